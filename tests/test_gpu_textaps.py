"""-m gpu: the opt-in texture-unit filter for the LIGHT segments' density taps (nrc_volume_set_tap_filter).  Control flow is untouched
(walks and vertex-to-vertex segments stay exact): every probe is bit-identical; radiance VALUES move by the hardware filter's 9-bit
weights -- stated bound: frame PSNR >= 50 dB against the oracle, >= 55 dB against the exact path."""
import numpy as np
import pytest

import common as Cm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def N():
    import neuralradiancecaching_b200 as N
    return N


@pytest.fixture()
def ctx(N):
    c = N.Context(0)
    yield c
    c.destroy()


@pytest.mark.parametrize("use_nn", [0, 1])
def test_texture_taps_keep_control_flow_and_stay_close(N, ctx, use_nn):
    from test_gpu_frame import build
    W, H = 96, 64
    cfg = Cm.scene_cfg(use_nn=use_nn, aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    orgba, ocnt, oqry = sc.render(W, H)
    frames, probes = [], []
    for tex in (False, True):
        r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
        vol.set_tap_filter(tex)
        r.FrameRender(); ctx.sync()
        frames.append(r.read_frame())
        probes.append(r.read_render_probes())
        r.Destroy()
    (c0, q0), (c1, q1) = probes
    np.testing.assert_array_equal(c0, c1)                          # RNG draws, lit vertices, walk steps, SDF steps
    np.testing.assert_array_equal(q0.view(np.uint32), q1.view(np.uint32))   # emitted queries
    np.testing.assert_array_equal(c1[..., 0], ocnt[..., 0])
    # background decisions (miss or T == 1.0 exactly -> env colour, fwd.frag:848-864) are control flow: identical pixel sets
    a0, a1 = frames[0][..., 3], frames[1][..., 3]
    np.testing.assert_array_equal(a0 == 1.0, a1 == 1.0)
    assert Cm.psnr(frames[1], frames[0]) >= 55.0, Cm.psnr(frames[1], frames[0])


def test_texture_taps_whole_frame_with_training(N, ctx):
    """train -> steps -> render with hardware-filtered light taps in BOTH passes: PSNR against the oracle's frame"""
    from test_gpu_frame import build
    W, H = 64, 48
    cfg = Cm.scene_cfg(aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    oframe = sc.frame(W, H, 8, 8)
    r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
    vol.set_tap_filter(True)
    r.Render(); ctx.sync()
    frame = r.read_frame()
    p = Cm.psnr(frame, oframe)
    assert p >= 50.0, p
    r.Destroy()
