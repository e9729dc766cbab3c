"""Deterministic synthetic stand-ins for the reference's missing data blobs (.MISSING_LARGE_BLOBS): the
125x85x153 density grid `data/cloud_sixteenth` and an HDR environment map.  Integer-hash based (no RNG library), so
every machine generates bit-identical arrays; SHA-256 of the default grid is recorded in BASELINE.md."""
import hashlib

import numpy as np


def _hash3(ix, iy, iz, seed):
    h = (ix.astype(np.uint64) * np.uint64(73856093)) ^ (iy.astype(np.uint64) * np.uint64(19349663)) ^ \
        (iz.astype(np.uint64) * np.uint64(83492791)) ^ np.uint64(seed * 2654435761 & 0xFFFFFFFF)
    h = (h ^ (h >> np.uint64(13))) * np.uint64(1274126177) & np.uint64(0xFFFFFFFF)
    h = h ^ (h >> np.uint64(16))
    return (h & np.uint64(0xFFFFFF)).astype(np.float64) / float(1 << 24)


def _value_noise(x, y, z, freq, seed):
    x, y, z = x * freq, y * freq, z * freq
    ix, iy, iz = np.floor(x).astype(np.int64), np.floor(y).astype(np.int64), np.floor(z).astype(np.int64)
    fx, fy, fz = x - ix, y - iy, z - iz
    sx, sy, sz = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy), fz * fz * (3 - 2 * fz)
    acc = 0.0
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                w = (sx if dx else 1 - sx) * (sy if dy else 1 - sy) * (sz if dz else 1 - sz)
                acc = acc + w * _hash3(ix + dx + 1000, iy + dy + 1000, iz + dz + 1000, seed)
    return acc


def cloud_density(nx=125, ny=85, nz=153, seed=7):
    """float32 [x][y][z] in [0, 1]: an fBm-eroded ellipsoid with empty space around it (about a third of the box is
    cloud), the shape ReadFileDensity3D(\"data/cloud_sixteenth\", 125, 85, 153) returns (src/main.cpp:123)."""
    x, y, z = np.meshgrid((np.arange(nx) + 0.5) / nx, (np.arange(ny) + 0.5) / ny, (np.arange(nz) + 0.5) / nz, indexing="ij")
    r = np.sqrt(((x - 0.5) / 0.42) ** 2 + ((y - 0.45) / 0.36) ** 2 + ((z - 0.5) / 0.44) ** 2)
    fbm = 0.0
    amp, freq = 0.5, 4.0
    for o in range(4):
        fbm = fbm + amp * _value_noise(x, y, z, freq, seed + o)
        amp, freq = amp * 0.5, freq * 2.0
    d = np.clip(1.0 - r + 1.1 * (fbm * 1.0667 - 0.5), 0.0, None) * 2.5
    return np.clip(d, 0.0, 1.0).astype(np.float32)


def refined_density(scale=4):
    """BASELINE config 5's "full-density cloud": the sixteenth-size grid refined `scale` times per axis by linear interpolation
    (500 x 340 x 612 at scale 4; the full-size file is not in the reference)."""
    from scipy import ndimage
    return np.clip(ndimage.zoom(cloud_density(), scale, order=1), 0.0, 1.0).astype(np.float32)


def sha256(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sky_envmap(width=256, height=128):
    """float32 RGBA [h][w][4]: horizon-to-zenith gradient plus a sun disc (HDR peak ~60)."""
    u = (np.arange(width) + 0.5) / width
    v = (np.arange(height) + 0.5) / height
    uu, vv = np.meshgrid(u, v)
    phi = (uu - 0.5) * 2 * np.pi
    theta = (vv - 0.5) * np.pi
    d = np.stack([np.cos(theta) * np.cos(phi), np.sin(theta), np.cos(theta) * np.sin(phi)], -1)
    sun = np.array([0.3, 0.8, -0.52])
    sun = sun / np.linalg.norm(sun)
    c = np.clip((d * sun).sum(-1), -1, 1)
    t = np.clip(d[..., 1] * 0.5 + 0.5, 0, 1)
    sky = np.stack([0.25 + 0.35 * (1 - t), 0.35 + 0.35 * (1 - t), 0.55 + 0.4 * t], -1)
    disc = np.exp((c - 1.0) * 900.0) * 60.0 + np.exp((c - 1.0) * 30.0) * 1.5
    rgb = sky + disc[..., None] * np.array([1.0, 0.92, 0.8])
    out = np.ones((height, width, 4), np.float32)
    out[..., :3] = rgb.astype(np.float32)
    return out


def init_params(seed, count, scale):
    """N(0,1)*scale float32 from a fixed numpy generator (parity runs set parameters explicitly)."""
    return (np.random.Generator(np.random.PCG64(seed)).standard_normal(count) * scale).astype(np.float32)


def mlp_init(seed=1):
    """W ~ N(0,1)*0.01, biases 0 (src/NeuralRadianceCache.cpp:629-646)."""
    p = np.zeros(20995, np.float32)
    p[:20672] = init_params(seed, 20672, 0.01)
    return p


def table_init(seed=2, L=16, T=16384, F=2):
    """tables ~ N(0,1)*0.1 (src/MRHE.cpp:127-135)."""
    return init_params(seed, L * T * F, 0.1)
