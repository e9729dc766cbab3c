"""-m gpu: parity at the sizes BASELINE.json's configs name, not only on toy frames: sampled rows of 1280x720 / 1920x1080 / 3840x2160
frames and the full 100x100 train grid against the oracle (control flow bit for bit), 2^22-entry hash tables (encode bit-exact, the
warp-aggregated global scatter of the backward pass), a whole frame with the as-shipped std140 resolution table, and the refined
500x340x612 grid of config 5."""
import numpy as np
import pytest

import common as Cm

pytestmark = pytest.mark.gpu

PSNR_MIN = 60.0


@pytest.fixture(scope="module")
def N():
    import neuralradiancecaching_b200 as N
    return N


@pytest.fixture(scope="module")
def ctx(N):
    c = N.Context(0)
    yield c
    c.destroy()


def _relmax(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) / max(float(np.max(np.abs(b))), 1e-30))


def _build(N, ctx, cfg, W, H, tw, th, density=None, res=None, T=16384, params=None, tables=None):
    vol = N.VolumeData(ctx, Cm.density() if density is None else density)
    vol.useNN, vol.showNonNN, vol.densityFactor, vol.g = cfg["use_nn"], cfg["show_non_nn"], cfg["density_factor"], cfg["g"]
    vol.random = cfg["random4"]
    env_rgba = Cm.envmap()
    env = N.HdrEnvMap(ctx, env_rgba.shape[1], env_rgba.shape[0], env_rgba)
    env.set_strengths(cfg["env_direct"], cfg["env_hpm"])
    dl = N.DirLight(Cm.DIR_ZENITH, Cm.DIR_AZIMUTH, (1, 1, 1), cfg["dir_strength"])
    pl = N.PointLight(cfg["pt_pos"], cfg["pt_color"], cfg["pt_strength"])
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    nrc.set_params(Cm.S.mlp_init(1) if params is None else params)
    m = N.MRHE(ctx, 0.01, 0.0, 16, T, 2, 16, 512, res, seed=0)
    m.set_tables(Cm.S.table_init(2, 16, T, 2) if tables is None else tables)
    r = N.NrcHpmRenderer(W, H, tw, th, cfg["cam"], vol, dl, pl, env, nrc, m)
    return r, nrc, m, vol, env


def _destroy(*objs):
    for o in objs:
        o.Destroy()


@pytest.mark.parametrize("W,H", [(1280, 720), (1920, 1080), (3840, 2160)])
def test_sampled_rows_of_the_baseline_frames(N, ctx, W, H):
    """Every 40th row (from row 17) of the frames configs 3, the headline and config 4 quote, rendered as a strided partition: RNG
    draws, lit vertices, walk steps, emitted queries and the queries' positions / directions bit for bit, pixels to 60 dB / 1e-3."""
    first, stride = 17, 40
    cfg = Cm.scene_cfg(use_nn=1, aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    r, nrc, m, vol, env = _build(N, ctx, cfg, W, H, 8, 8)
    r.SetPartitionStrided(first, stride, 0, 1)
    r.FrameRender()
    ctx.sync()
    cnt, qry = r.read_render_probes()
    frame = r.read_frame()
    rows = list(range(first, H, stride))
    assert len(rows) >= 18
    lit = 0
    for y in rows:
        orgba, ocnt, oqry = sc.render(W, H, y, y + 1)
        np.testing.assert_array_equal(cnt[y, :, 0], ocnt[y, :, 0], err_msg="draws, row %d" % y)
        np.testing.assert_array_equal(cnt[y, :, 1], ocnt[y, :, 1], err_msg="lit vertices, row %d" % y)
        np.testing.assert_array_equal(cnt[y, :, 2], ocnt[y, :, 3], err_msg="walk steps, row %d" % y)
        np.testing.assert_array_equal(cnt[y, :, 3], oqry[y, :, 7].astype(np.uint32), err_msg="queries, row %d" % y)
        np.testing.assert_array_equal(qry[y, :, :6].view(np.uint32), oqry[y, :, :6].view(np.uint32), err_msg="query bits, row %d" % y)
        assert Cm.psnr(frame[y], orgba[y], peak=max(float(np.abs(orgba[y]).max()), 1.0)) >= PSNR_MIN, y
        lit += int(ocnt[y, :, 1].sum())
    assert lit > 1000      # the sampled rows do cross the cloud
    _destroy(r, nrc, m, vol, env)


def test_full_train_grid_parity(N, ctx):
    """The bench's train pass: all 100 x 100 invocations (nrc-train.comp main, NrcHpmRenderer.cpp:604) -- counters and records bit for
    bit, targets 1e-4, accumulated gradients 1e-3 -- on the 16:9 camera of the 1080p frame."""
    tw = th = 100
    cfg = Cm.scene_cfg(aspect=1920 / 1080)
    sc = Cm.oracle_scene(cfg)
    sc.set_train_size(tw, th)
    orec, ocnt = sc.train_trace(tw, th)
    r, nrc, m, vol, env = _build(N, ctx, cfg, 64, 36, tw, th)
    r.FrameTrain()
    ctx.sync()
    rec, cnt = r.read_train_probes()
    np.testing.assert_array_equal(cnt[:, 0], ocnt[:, 0])
    np.testing.assert_array_equal(cnt[:, 1], ocnt[:, 1])
    np.testing.assert_array_equal(cnt[:, 2], ocnt[:, 3])
    np.testing.assert_array_equal(rec[:, :6].view(np.uint32), orec[:, :6].view(np.uint32))
    assert _relmax(rec[:, 6:], orec[:, 6:]) < 1e-4
    sc.train_accumulate(orec[:, 0:3], orec[:, 3:6], orec[:, 6:9])
    assert _relmax(nrc.get_params()[1], sc.get_mlp()[1]) < 1e-3
    assert _relmax(m.get_tables()[1][:2 * 16384], sc.get_mrhe()[1][:2 * 16384]) < 1e-3
    _destroy(r, nrc, m, vol, env)


def test_hash_tables_of_config5(N, ctx):
    """T = 2^22 entries per level (512 MiB of tables): hash indices and features bit for bit; inference 1e-3; the backward pass, whose
    scatter no longer fits shared memory (warp-aggregated global atomics, csrc/mrhe_scatter.cuh), 1e-3 on the touched head."""
    T = 1 << 22
    tables = Cm.S.table_init(2, 16, T, 2)
    params = Cm.S.mlp_init(1)
    params[20672:] = Cm.S.init_params(7, 323, 0.01)
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=params, tables=tables, T=T)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    nrc.set_params(params)
    m = N.MRHE(ctx, 0.01, 0.0, 16, T, 2, 16, 512, None, seed=0)
    m.set_tables(tables)
    del tables
    pos, d = Cm.rand_queries(1 << 14, seed=41)
    pos[:4] = [[0, 0, 0], [31.25, 21.25, 38.25], [-31.25, -21.25, -38.25], [40, 40, 40]]
    gf, gi = N.encode(m, pos)
    of, oi, _ = sc.encode_pos(pos)
    assert int(oi.max()) >= (1 << 21)                                   # indices beyond the small tables' range
    np.testing.assert_array_equal(gi, oi)
    np.testing.assert_array_equal(gf.view(np.uint32), of.view(np.uint32))
    assert _relmax(N.infer(nrc, m, pos, d), sc.infer(pos, d)) < 1e-3
    n = 4096
    target = (np.random.Generator(np.random.PCG64(4)).random((n, 3)) * 3.0).astype(np.float32)
    sc.set_loss_scale(1.0 / n, 1.0)
    sc.train_accumulate(pos[:n], d[:n], target)
    N.train_accumulate(nrc, m, pos[:n], d[:n], target, 1.0 / n)
    assert _relmax(nrc.get_params()[1], sc.get_mlp()[1]) < 1e-3
    gt, ot = m.get_tables()[1], sc.get_mrhe()[1]
    assert np.count_nonzero(gt[2 * T:]) == 0                              # level 0's slice only
    assert np.count_nonzero(ot[:2 * T]) > 1000
    assert _relmax(gt[:2 * T], ot[:2 * T]) < 1e-3
    assert m.GetHashTableSize() == 16 * T * 2 * 4
    _destroy(nrc, m)


def test_whole_frames_with_the_as_shipped_resolution_table(N, ctx):
    """Render() frames with the resolution table the shipped SPIR-V actually reads (std140: res[1], res[5], res[9], res[13], then
    zeros) against the oracle in the same mode: frames 60 dB, trained weights and tables 1e-4."""
    W, H, tw, th = 96, 64, 16, 16
    cfg = Cm.scene_cfg(aspect=W / H)
    sc = Cm.oracle_scene(cfg, res=Cm.RES_SHIPPED)
    r, nrc, m, vol, env = _build(N, ctx, cfg, W, H, tw, th, res=Cm.RES_SHIPPED)
    for f, rnd in enumerate([(0.37, 0.61, 0.13, 0.89), (0.11, 0.93, 0.5, 0.2)]):
        sc.set_volume_params(rnd, 1, 0, 0.4, 0.7, 1, 1)
        sc.reset_stats()
        oframe = sc.frame(W, H, tw, th)
        vol.random = np.asarray(rnd, np.float32)
        nrc.ResetStats()
        r.Render()
        ctx.sync()
        assert Cm.psnr(r.read_frame(), oframe) >= PSNR_MIN, f
        assert abs(nrc.GetStats().mseLoss - sc.mse_loss()) <= 2e-3 * abs(sc.mse_loss()) + 1e-12
    assert _relmax(nrc.get_params()[0], sc.get_mlp()[0]) < 1e-4
    assert _relmax(m.get_tables()[0], sc.get_mrhe()[0]) < 1e-4
    _destroy(r, nrc, m, vol, env)


def test_refined_grid_of_config5(N, ctx):
    """The 500 x 340 x 612 grid (838 MB of bricks, no longer L2-resident): getDensity zero-ness exactly and values 1e-5 against the
    oracle's sampler; a small render pass through it bit for bit in its control flow."""
    dens = Cm.S.refined_density(4)
    assert dens.shape == (500, 340, 612)
    W, H = 96, 64
    cfg = Cm.scene_cfg(use_nn=1, aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    sc.set_volume(dens)
    r, nrc, m, vol, env = _build(N, ctx, cfg, W, H, 8, 8, density=dens)
    vol.push()
    pos, _ = Cm.rand_queries(1 << 14, seed=9)
    pos *= 1.05
    got, want = vol.sample(pos), sc.density(pos)
    np.testing.assert_array_equal(got > 0, want > 0)
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-7)
    orgba, ocnt, oqry = sc.render(W, H)
    r.FrameRender()
    ctx.sync()
    cnt, qry = r.read_render_probes()
    np.testing.assert_array_equal(cnt[..., 0], ocnt[..., 0])
    np.testing.assert_array_equal(cnt[..., 1], ocnt[..., 1])
    np.testing.assert_array_equal(cnt[..., 2], ocnt[..., 3])
    np.testing.assert_array_equal(qry[..., :6].view(np.uint32), oqry[..., :6].view(np.uint32))
    assert Cm.psnr(r.read_frame(), orgba) >= PSNR_MIN
    _destroy(r, nrc, m, vol, env)
