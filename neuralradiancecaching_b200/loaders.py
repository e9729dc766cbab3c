"""The reference's data loaders (src/read_file.cpp) for the formats either side of the hot path.

ReadFileDensity3D  raw float32 grid, index = x*Y*Z + y*Z + z                       (read_file.cpp:50-78)
ReadFileHdr4f      Radiance RGBE (.hdr) -> float RGBA, flipped vertically like
                   stbi_set_flip_vertically_on_load(true) + stbi_loadf(..., 4)      (read_file.cpp:80-121)
Hdr4fToCdf         on the device: objects.hdr4f_to_cdf                              (read_file.cpp:141-206)

stb_image is not on this image; the RGBE decode below restates its published algorithm (stb_image.h, stbi__hdr_load /
stbi__hdr_convert: value = mantissa_byte * 2^(e - 136), alpha = 1, new-style RLE scanlines and flat files)."""
import numpy as np


def ReadFileDensity3D(file_name, x_size, y_size, z_size):
    raw = np.fromfile(file_name, dtype="<f4", count=x_size * y_size * z_size)
    if raw.size != x_size * y_size * z_size:
        raise RuntimeError("Failed to read %d floats from %s" % (x_size * y_size * z_size, file_name))
    return raw.reshape(x_size, y_size, z_size)


def WriteFileDensity3D(file_name, density_xyz):
    np.ascontiguousarray(density_xyz, dtype="<f4").tofile(file_name)


def _rgbe_to_float(rgbe):
    """rgbe: uint8 [..., 4] -> float32 [..., 4] (alpha 1), stbi__hdr_convert"""
    e = rgbe[..., 3].astype(np.int32)
    scale = np.where(e != 0, np.ldexp(np.float32(1.0), e - 136), np.float32(0.0)).astype(np.float32)
    out = np.ones(rgbe.shape[:-1] + (4,), np.float32)
    out[..., :3] = rgbe[..., :3].astype(np.float32) * scale[..., None]
    return out


def decode_hdr(data):
    """bytes of a Radiance .hdr file -> (width, height, uint8 RGBE [h][w][4]) in file (top-down) order"""
    pos = 0

    def line():
        nonlocal pos
        end = data.index(b"\n", pos)
        s = data[pos:end]
        pos = end + 1
        return s

    first = line()
    if first not in (b"#?RADIANCE", b"#?RGBE"):
        raise RuntimeError("not a Radiance HDR file")
    valid = False
    while True:
        s = line()
        if s == b"":
            break
        if s == b"FORMAT=32-bit_rle_rgbe":
            valid = True
    if not valid:
        raise RuntimeError("unsupported HDR format")
    tok = line().split()
    if len(tok) != 4 or tok[0] != b"-Y" or tok[2] != b"+X":
        raise RuntimeError("unsupported HDR data layout")
    height, width = int(tok[1]), int(tok[3])
    buf = np.frombuffer(data, dtype=np.uint8)
    img = np.zeros((height, width, 4), np.uint8)
    if width < 8 or width >= 32768:   # flat data
        img[:] = buf[pos:pos + width * height * 4].reshape(height, width, 4)
        return width, height, img
    for y in range(height):
        c1, c2, ln = int(buf[pos]), int(buf[pos + 1]), int(buf[pos + 2])
        if c1 != 2 or c2 != 2 or (ln & 0x80):
            # not run-length encoded: the whole remaining image is flat (stb's fallback on the first scanline)
            if y != 0:
                raise RuntimeError("corrupt HDR: mixed flat and RLE scanlines")
            img[:] = buf[pos:pos + width * height * 4].reshape(height, width, 4)
            return width, height, img
        if ((ln << 8) | int(buf[pos + 3])) != width:
            raise RuntimeError("corrupt HDR: invalid decoded scanline length")
        pos += 4
        for k in range(4):
            i = 0
            while i < width:
                count = int(buf[pos]); pos += 1
                if count > 128:
                    count -= 128
                    if count == 0 or i + count > width:
                        raise RuntimeError("corrupt HDR run")
                    img[y, i:i + count, k] = buf[pos]; pos += 1
                else:
                    if count == 0 or i + count > width:
                        raise RuntimeError("corrupt HDR run")
                    img[y, i:i + count, k] = buf[pos:pos + count]; pos += count
                i += count
    return width, height, img


def ReadFileHdr4f(file_name):
    """-> (hdr4f float32 [h][w][4] with row 0 = BOTTOM row of the picture, width, height)"""
    with open(file_name, "rb") as f:
        w, h, rgbe = decode_hdr(f.read())
    return np.ascontiguousarray(_rgbe_to_float(rgbe)[::-1]), w, h


def encode_hdr(rgbe):
    """uint8 RGBE [h][w][4] (top-down) -> bytes, new-style RLE; the inverse of decode_hdr (used by the tests and to write
    fixtures)."""
    h, w = rgbe.shape[:2]
    out = bytearray(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w))
    if w < 8 or w >= 32768:
        out += rgbe.tobytes()
        return bytes(out)
    for y in range(h):
        out += bytes([2, 2, (w >> 8) & 0xff, w & 0xff])
        for k in range(4):
            row = rgbe[y, :, k]
            i = 0
            while i < w:
                run = 1
                while i + run < w and run < 127 and row[i + run] == row[i]:
                    run += 1
                if run >= 3:
                    out += bytes([128 + run, int(row[i])])
                    i += run
                else:
                    j = i
                    while j < w and j - i < 128:
                        if j + 2 < w and row[j] == row[j + 1] == row[j + 2]:
                            break
                        j += 1
                    j = max(j, i + 1)
                    out += bytes([j - i]) + row[i:j].tobytes()
                    i = j
    return bytes(out)


def float_to_rgbe(rgb):
    """float32 [..., 3] -> uint8 RGBE (for fixtures)"""
    rgb = np.asarray(rgb, np.float32)
    m = rgb.max(axis=-1)
    mant, e = np.frexp(m)
    scale = np.where(m > 1e-32, mant * 256.0 / np.maximum(m, 1e-32), 0.0)
    out = np.zeros(rgb.shape[:-1] + (4,), np.uint8)
    out[..., :3] = np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8)
    out[..., 3] = np.where(m > 1e-32, e + 128, 0).astype(np.uint8)
    return out
