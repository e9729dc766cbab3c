"""Loaders either side of the hot path: raw density grid and Radiance .hdr."""
import os

import numpy as np
import pytest

from neuralradiancecaching_b200 import loaders as Ld
from neuralradiancecaching_b200 import scene as S


def test_density_roundtrip_and_order(tmp_path):
    d = np.arange(3 * 4 * 5, dtype=np.float32).reshape(3, 4, 5)
    p = str(tmp_path / "grid")
    Ld.WriteFileDensity3D(p, d)
    raw = np.fromfile(p, "<f4")
    assert raw[1 * 4 * 5 + 2 * 5 + 3] == d[1, 2, 3]          # index = x*Y*Z + y*Z + z (read_file.cpp:61-74)
    np.testing.assert_array_equal(Ld.ReadFileDensity3D(p, 3, 4, 5), d)
    with pytest.raises(RuntimeError):
        Ld.ReadFileDensity3D(p, 3, 4, 6)


@pytest.mark.parametrize("w,h", [(64, 16), (5, 7), (300, 3)])
def test_hdr_rle_decode_matches_rgbe_semantics(tmp_path, w, h):
    env = S.sky_envmap(w, h)[..., :3]
    env[:, : w // 3] = env[:1, :1]                              # long constant runs exercise the RLE branch
    rgbe = Ld.float_to_rgbe(env)
    p = str(tmp_path / "e.hdr")
    open(p, "wb").write(Ld.encode_hdr(rgbe))
    got, gw, gh = Ld.ReadFileHdr4f(p)
    assert (gw, gh) == (w, h) and got.shape == (h, w, 4)
    # stbi__hdr_convert: mantissa * 2^(e-136), alpha 1; vertical flip (stbi_set_flip_vertically_on_load)
    want = np.ones((h, w, 4), np.float32)
    want[..., :3] = rgbe[..., :3].astype(np.float32) * np.ldexp(np.float32(1), rgbe[..., 3].astype(np.int32) - 136)[..., None]
    np.testing.assert_array_equal(got, want[::-1])
    assert np.abs(got[::-1, :, :3] - env).max() <= env.max() / 128.0   # RGBE quantisation of the source


def test_hdr_rejects_garbage(tmp_path):
    p = str(tmp_path / "x.hdr")
    open(p, "wb").write(b"P6\n1 1\n255\n\0\0\0")
    with pytest.raises(RuntimeError):
        Ld.ReadFileHdr4f(p)


def test_reference_photostudio_against_opencv():
    """Only where the reference checkout and OpenCV exist (this container): decode the bundled map with both."""
    path = "/root/reference/data/image/photostudio.hdr"
    cv2 = pytest.importorskip("cv2")
    if not os.path.exists(path):
        pytest.skip("reference data not present")
    got, w, h = Ld.ReadFileHdr4f(path)
    ref = cv2.imread(path, cv2.IMREAD_UNCHANGED)                # BGR float32, top-down
    assert ref is not None and ref.shape[:2] == (h, w)
    np.testing.assert_allclose(got[::-1, :, :3], ref[..., ::-1], rtol=1e-6, atol=0)


def test_scene_hash():
    assert S.sha256(S.cloud_density()) == "dad76e8100b2873e18dd8b1154ff24e11e55020936673300776d1383428014ab"
