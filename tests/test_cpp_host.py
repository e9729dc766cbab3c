"""The C++ host side either side of the hot path (include/nrc_host.hpp): raw density reader, Radiance .hdr reader, frame writers and
the NrcConfig option parser, through the CPU-only examples/host_tool.cpp.  The readers are checked against the Python loaders (which
are checked against OpenCV) and, where the reference checkout exists, against OpenCV directly on the reference's photostudio.hdr."""
import os
import subprocess

import numpy as np
import pytest

from neuralradiancecaching_b200 import loaders as Ld
from neuralradiancecaching_b200 import scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "_build", "host_tool")


@pytest.fixture(scope="module")
def tool():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "host_tool.cpp"), "-o", EXE])
    return EXE


def _run(tool, *args):
    return subprocess.run([tool, *args], capture_output=True, text=True, timeout=120)


def test_density_reader_order_and_short_file(tool, tmp_path):
    d = np.arange(3 * 4 * 5, dtype=np.float32).reshape(3, 4, 5)
    src, dst = str(tmp_path / "grid"), str(tmp_path / "back")
    Ld.WriteFileDensity3D(src, d)
    p = _run(tool, "density", src, "3", "4", "5", dst)
    assert p.returncode == 0 and p.stdout.strip() == "60"
    np.testing.assert_array_equal(np.fromfile(dst, "<f4").reshape(3, 4, 5), d)        # index = x*Y*Z + y*Z + z (read_file.cpp:61-74)
    p = _run(tool, "density", src, "3", "4", "6", dst)
    assert p.returncode == 1 and "Failed to read 72 floats" in p.stderr
    p = _run(tool, "density", str(tmp_path / "missing"), "1", "1", "1", dst)
    assert p.returncode == 1 and "Failed to open" in p.stderr


@pytest.mark.parametrize("w,h", [(64, 16), (5, 7), (300, 3)])
def test_hdr_reader_equals_the_python_loader(tool, tmp_path, w, h):
    env = S.sky_envmap(w, h)[..., :3]
    env[:, : w // 3] = env[:1, :1]                              # long constant runs exercise the RLE branch
    src, dst = str(tmp_path / "e.hdr"), str(tmp_path / "e.raw")
    open(src, "wb").write(Ld.encode_hdr(Ld.float_to_rgbe(env)))
    p = _run(tool, "hdr", src, dst)
    assert p.returncode == 0 and p.stdout.split() == [str(w), str(h)]
    want, _, _ = Ld.ReadFileHdr4f(src)
    np.testing.assert_array_equal(np.fromfile(dst, "<f4").reshape(h, w, 4), want)      # bit for bit, flipped, alpha 1


def test_hdr_reader_rejects_garbage(tool, tmp_path):
    src = str(tmp_path / "x.hdr")
    open(src, "wb").write(b"P6\n1 1\n255\n\0\0\0")
    p = _run(tool, "hdr", src, str(tmp_path / "o"))
    assert p.returncode == 1 and "not a Radiance HDR file" in p.stderr
    open(src, "wb").write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 4 +X 16\n\x02\x02\x00\x10\x90")   # truncated run
    p = _run(tool, "hdr", src, str(tmp_path / "o"))
    assert p.returncode == 1 and "corrupt HDR" in p.stderr


def test_reference_photostudio_against_opencv(tool, tmp_path):
    """Only where the reference checkout and OpenCV exist (this container): the C++ reader on the map the reference bundles."""
    path = "/root/reference/data/image/photostudio.hdr"
    cv2 = pytest.importorskip("cv2")
    if not os.path.exists(path):
        pytest.skip("reference data not present")
    dst = str(tmp_path / "p.raw")
    p = _run(tool, "hdr", path, dst)
    assert p.returncode == 0
    w, h = (int(x) for x in p.stdout.split())
    ref = cv2.imread(path, cv2.IMREAD_UNCHANGED)                # BGR float32, top-down
    assert ref is not None and ref.shape[:2] == (h, w)
    got = np.fromfile(dst, "<f4").reshape(h, w, 4)
    np.testing.assert_allclose(got[::-1, :, :3], ref[..., ::-1], rtol=1e-6, atol=0)
    assert (got[..., 3] == 1.0).all()


def test_frame_writers(tool, tmp_path):
    pfm, ppm = str(tmp_path / "f.pfm"), str(tmp_path / "f.ppm")
    assert _run(tool, "pfm", pfm).returncode == 0 and _run(tool, "ppm", ppm).returncode == 0
    data = open(pfm, "rb").read()
    head = b"PF\n3 2\n-1.0\n"
    assert data.startswith(head)
    px = np.frombuffer(data[len(head):], "<f4").reshape(2, 3, 3)
    np.testing.assert_array_equal(px[0, 0], [7, 8, 9])          # PFM rows run bottom-up: the first stored row is the frame's last
    np.testing.assert_array_equal(px[1, 0], [0, 0.5, 1])
    data = open(ppm, "rb").read()
    assert data.startswith(b"P6\n3 2\n255\n") and list(data[11:17]) == [1, 2, 3, 4, 5, 6]   # BGRA8 -> RGB


def test_config_defaults_are_the_reference_literals(tool):
    p = _run(tool, "config")
    kv = dict(l.split("=", 1) for l in p.stdout.strip().splitlines())
    assert kv["width"] == "800" and kv["height"] == "800" and kv["train"] == "100x100"           # main.cpp:112-113, 161
    assert kv["grid"] == "125,85,153" and kv["cam"] == "0,0,-64" and kv["fov"] == "60"           # main.cpp:123, 142-149
    assert kv["lr"] == "0.001,0,0.5" and kv["mrhe"] == "0.01,0,16,16384,16,512"                  # main.cpp:156-157, MRHE.cpp:83-87
    assert kv["g"] == "0.7" and kv["densityFactor"] == "0.4" and kv["spp"] == "1,1" and kv["env"] == "1,8"
    assert kv["resolutions"] == "intended" and kv["precision"] == "bf16" and kv["gpus"] == "1"


def test_config_parser_options_and_errors(tool):
    p = _run(tool, "config", "--width=1280", "--height=720", "--spp=2", "--table-size=4194304", "--point-pos=1,2,3", "--point-strength=0.5",
             "--fixed", "--adam=0", "--resolutions=as_shipped_std140", "--precision=fp32", "--gpus=8", "--rank=3", "--out-pfm=x.pfm",
             "--density=cloud_sixteenth", "--envmap=photostudio.hdr")
    assert p.returncode == 0, p.stderr
    kv = dict(l.split("=", 1) for l in p.stdout.strip().splitlines())
    assert kv["width"] == "1280" and kv["height"] == "720" and kv["spp"] == "2,2" and kv["mrhe"] == "0.01,0,16,4194304,16,512"
    assert kv["point"] == "1,2,3,0.5" and kv["fixed"] == "1" and kv["adam"] == "0" and kv["resolutions"] == "as_shipped_std140"
    assert kv["precision"] == "fp32" and kv["gpus"] == "8" and kv["rank"] == "3" and kv["outPfm"] == "x.pfm"
    assert kv["density"] == "cloud_sixteenth" and kv["envmap"] == "photostudio.hdr"
    for bad, msg in (("--bogus=1", "unknown option"), ("--width=abc", "bad value"), ("--width", "bad value"), ("frames", "unexpected argument"),
                     ("--point-pos=1,2", "bad value"), ("--precision=fp64", "--precision must be"), ("--rank=2", "--rank must be"),
                     ("--levels=17", "--levels must be")):
        p = _run(tool, "config", bad)
        assert p.returncode == 1 and msg in p.stderr, (bad, p.stderr)
