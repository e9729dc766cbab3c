"""-m gpu: the tracer and the whole frame (train -> steps -> render) against the oracle.
Control flow (RNG draw counts, lit vertices, walk steps, emitted queries, query/record positions and directions) is
compared BIT FOR BIT; radiance values within fp32 tolerance; frames by PSNR (stated: >= 60 dB on the float frame)."""
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


def build(N, ctx, cfg, W, H, tw, th, params=None, tables=None):
    vol = N.VolumeData(ctx, Cm.density())
    vol.useNN, vol.showNonNN, vol.densityFactor, vol.g = cfg["use_nn"], cfg["show_non_nn"], cfg["density_factor"], cfg["g"]
    vol.random = cfg["random4"]
    env_rgba = Cm.envmap()
    env = N.HdrEnvMap(ctx, env_rgba.shape[1], env_rgba.shape[0], env_rgba)
    env.set_strengths(cfg["env_direct"], cfg["env_hpm"])
    dl = N.DirLight(Cm.DIR_ZENITH, Cm.DIR_AZIMUTH, (1, 1, 1), cfg["dir_strength"])
    pl = N.PointLight(cfg["pt_pos"], cfg["pt_color"], cfg["pt_strength"])
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    nrc.set_params(Cm.S.mlp_init(1) if params is None else params)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0)
    m.set_tables(Cm.S.table_init() if tables is None else tables)
    r = N.NrcHpmRenderer(W, H, tw, th, cfg["cam"], vol, dl, pl, env, nrc, m)
    return r, nrc, m, vol


def _relmax(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) / max(float(np.max(np.abs(b))), 1e-30))


@pytest.mark.parametrize("use_nn", [1, 0])
def test_render_pass_parity(N, ctx, use_nn):
    W, H = 96, 64
    cfg = Cm.scene_cfg(use_nn=use_nn, aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    orgba, ocnt, oqry = sc.render(W, H)
    r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
    r.FrameRender()
    ctx.sync()
    cnt, qry = r.read_render_probes()
    np.testing.assert_array_equal(cnt[..., 0], ocnt[..., 0])              # RNG draws
    np.testing.assert_array_equal(cnt[..., 1], ocnt[..., 1])              # lit vertices
    np.testing.assert_array_equal(cnt[..., 2], ocnt[..., 3])              # walk steps
    np.testing.assert_array_equal(cnt[..., 3], oqry[..., 7].astype(np.uint32))   # query emitted
    np.testing.assert_array_equal(qry[..., :6].view(np.uint32), oqry[..., :6].view(np.uint32))
    np.testing.assert_allclose(qry[..., 6], oqry[..., 6], rtol=2e-4, atol=1e-7)   # T * dirPhase
    frame = r.read_frame()
    assert Cm.psnr(frame, orgba) >= PSNR_MIN
    assert _relmax(frame, orgba) < 1e-3
    st = r.stats()
    assert st["infer_queries"] == int(oqry[..., 7].sum())
    # density taps of the light kernel == the oracle's taps minus one per walk step
    assert st["density_taps"] == int(ocnt[..., 2].sum()) - int(ocnt[..., 3].sum())
    # BGRA8 frame = clamp + round of the float frame, B,G,R,A byte order (VK_FORMAT_B8G8R8A8_UNORM)
    f8 = r.read_frame(N._lib.FRAME_BGRA8)
    want8 = np.rint(np.clip(frame[..., [2, 1, 0, 3]], 0, 1) * 255).astype(np.uint8)
    assert np.abs(f8.astype(int) - want8.astype(int)).max() <= 1
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_split_render_walk_is_the_single_launch_walk(N, ctx, monkeypatch):
    """NRC_RWALK_SPLIT: the head launch walks k steps, the tail launch finishes the compacted survivors.  Every split point gives the
    counters, queries and frame of the single-launch walk (k = 0), bit for bit; the default (10) is what the other tests run with."""
    W, H = 160, 96
    cfg = Cm.scene_cfg(use_nn=1, aspect=W / H)
    got = {}
    for k in (0, 1, 10, 31):
        monkeypatch.setenv("NRC_RWALK_SPLIT", str(k))
        r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
        r.FrameRender()
        ctx.sync()
        cnt, qry = r.read_render_probes()
        got[k] = (cnt.copy(), qry.copy(), r.read_frame().copy(), r.stats())
        r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()
    monkeypatch.delenv("NRC_RWALK_SPLIT")
    cnt0, qry0, frame0, st0 = got[0]
    assert cnt0[..., 2].max() == 32 and (cnt0[..., 2] < 8).any()          # both kinds of path are in the frame
    for k in (1, 10, 31):
        cnt, qry, frame, st = got[k]
        np.testing.assert_array_equal(cnt, cnt0)
        np.testing.assert_array_equal(qry.view(np.uint32), qry0.view(np.uint32))
        np.testing.assert_array_equal(frame.view(np.uint32), frame0.view(np.uint32))
        assert st["density_taps"] == st0["density_taps"] and st["infer_queries"] == st0["infer_queries"]
    # against the oracle as well, at the default split
    sc = Cm.oracle_scene(cfg)
    orgba, ocnt, oqry = sc.render(W, H)
    np.testing.assert_array_equal(got[10][0][..., 0], ocnt[..., 0])
    np.testing.assert_array_equal(got[10][0][..., 2], ocnt[..., 3])
    assert Cm.psnr(got[10][2], orgba) >= PSNR_MIN


def test_train_pass_parity(N, ctx):
    tw, th = 20, 12
    cfg = Cm.scene_cfg()
    sc = Cm.oracle_scene(cfg)
    sc.set_train_size(tw, th)
    orec, ocnt = sc.train_trace(tw, th)
    r, nrc, m, vol = build(N, ctx, cfg, 32, 32, tw, th)
    r.FrameTrain()
    ctx.sync()
    rec, cnt = r.read_train_probes()
    np.testing.assert_array_equal(cnt[:, 0], ocnt[:, 0])
    np.testing.assert_array_equal(cnt[:, 1], ocnt[:, 1])
    np.testing.assert_array_equal(cnt[:, 2], ocnt[:, 3])
    np.testing.assert_array_equal(rec[:, :6].view(np.uint32), orec[:, :6].view(np.uint32))
    assert _relmax(rec[:, 6:], orec[:, 6:]) < 1e-4
    # gradients accumulated by the same call == oracle Backprop over the oracle's records
    sc.train_accumulate(orec[:, 0:3], orec[:, 3:6], orec[:, 6:9])
    _, od, _ = sc.get_mlp()
    _, gd, _ = nrc.get_params()
    assert _relmax(gd, od) < 1e-3
    assert abs(nrc.GetStats().mseLoss - sc.mse_loss()) <= 1e-3 * abs(sc.mse_loss()) + 1e-12
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_whole_frames_parity_and_partition(N, ctx):
    """Three Render() frames (train -> nrc-step -> mrhe-step -> render, new `random` each frame) against orc_frame; then the
    same frame rendered as two row bands + two train shards must be bit-identical to the single-partition frame."""
    W, H, tw, th = 64, 48, 16, 16
    cfg = Cm.scene_cfg(aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    r, nrc, m, vol = build(N, ctx, cfg, W, H, tw, th)
    rnd = [(0.37, 0.61, 0.13, 0.89), (0.11, 0.93, 0.5, 0.2), (0.71, 0.29, 0.4, 0.6)]
    for f in range(3):
        sc.set_volume_params(rnd[f], 1, 0, 0.4, 0.7, 1, 1)
        sc.reset_stats()
        oframe = sc.frame(W, H, tw, th)
        vol.random = np.asarray(rnd[f], np.float32)
        nrc.ResetStats()
        r.Render()
        ctx.sync()
        frame = r.read_frame()
        assert Cm.psnr(frame, oframe) >= PSNR_MIN, f
        assert abs(nrc.GetStats().mseLoss - sc.mse_loss()) <= 2e-3 * abs(sc.mse_loss()) + 1e-12
    assert _relmax(nrc.get_params()[0], sc.get_mlp()[0]) < 1e-4
    assert _relmax(m.get_tables()[0], sc.get_mrhe()[0]) < 1e-4
    # partition invariance of the render pass (weights fixed now)
    r.FrameRender(); ctx.sync()
    full = r.read_frame().copy()
    r.SetPartition(0, H // 2, 0, tw * th // 2); r.FrameRender(); ctx.sync()
    top = r.read_frame()[:H // 2].copy()
    r.SetPartition(H // 2, H, tw * th // 2, tw * th); r.FrameRender(); ctx.sync()
    bottom = r.read_frame()[H // 2:].copy()
    np.testing.assert_array_equal(np.concatenate([top, bottom]).view(np.uint32), full.view(np.uint32))
    # interleaved (strided) partition: rank k of 3 renders rows k, k+3, ...; compacted read-back, union == full frame
    union = np.zeros_like(full)
    for k in range(3):
        r.SetPartitionStrided(k, 3, k, 3); r.FrameRender(); ctx.sync()
        nrows = len(range(k, H, 3))
        assert r.partition_bytes(N._lib.FRAME_RGBA32F) == nrows * W * 16
        part = np.zeros((nrows, W, 4), np.float32)
        r.read_frame_partition(part, N._lib.FRAME_RGBA32F)
        union[k::3] = part
    np.testing.assert_array_equal(union.view(np.uint32), full.view(np.uint32))
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_ragged_sizes_parity(N, ctx):
    """A frame whose width and height are no multiple of the render walk's 8 x 4 warp tile, a train grid that is no multiple of a warp,
    and a strided partition whose row count is no multiple of 4: counters, queries and training records bit-exact against the oracle
    (the ragged edges of every launch: partial tiles, partial warps, partial 128-row MLP tiles)."""
    W, H, tw, th = 61, 37, 7, 5
    cfg = Cm.scene_cfg(use_nn=1, aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    sc.set_train_size(tw, th)
    orgba, ocnt, oqry = sc.render(W, H)
    orec, otc = sc.train_trace(tw, th)
    r, nrc, m, vol = build(N, ctx, cfg, W, H, tw, th)
    r.FrameRender()
    ctx.sync()
    cnt, qry = r.read_render_probes()
    np.testing.assert_array_equal(cnt[..., 0], ocnt[..., 0])
    np.testing.assert_array_equal(cnt[..., 1], ocnt[..., 1])
    np.testing.assert_array_equal(cnt[..., 2], ocnt[..., 3])
    np.testing.assert_array_equal(cnt[..., 3], oqry[..., 7].astype(np.uint32))
    np.testing.assert_array_equal(qry[..., :6].view(np.uint32), oqry[..., :6].view(np.uint32))
    full = r.read_frame().copy()
    assert Cm.psnr(full, orgba) >= PSNR_MIN
    r.FrameTrain()
    ctx.sync()
    rec, tc = r.read_train_probes()
    np.testing.assert_array_equal(tc[:, 0], otc[:, 0])
    np.testing.assert_array_equal(tc[:, 1], otc[:, 1])
    np.testing.assert_array_equal(tc[:, 2], otc[:, 3])
    np.testing.assert_array_equal(rec[:, :6].view(np.uint32), orec[:, :6].view(np.uint32))
    # rank 1 of 3: rows 1, 4, ... (12 of 37 rows) and samples 1, 4, ... (12 of 35)
    union = np.zeros_like(full)
    for k in range(3):
        r.SetPartitionStrided(k, 3, k, 3); r.FrameRender(); ctx.sync()
        nrows = len(range(k, H, 3))
        part = np.zeros((nrows, W, 4), np.float32)
        r.read_frame_partition(part, N._lib.FRAME_RGBA32F)
        union[k::3] = part
    np.testing.assert_array_equal(union.view(np.uint32), full.view(np.uint32))
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_resize_and_spp(N, ctx):
    cfg = Cm.scene_cfg(use_nn=0)
    r, nrc, m, vol = build(N, ctx, cfg, 32, 24, 4, 4)
    r.ResizeFrame(48, 40)
    cfg2 = Cm.scene_cfg(use_nn=0, aspect=48 / 40)
    r.camera = cfg2["cam"]
    vol.noNnSpp = 2
    sc = Cm.oracle_scene(cfg2)
    sc.set_volume_params(cfg2["random4"], 0, 0, 0.4, 0.7, 2, 1)
    orgba, _, _ = sc.render(48, 40)
    r.FrameRender(); ctx.sync()
    assert r.GetImageDataSize() == 48 * 40 * 4
    assert Cm.psnr(r.read_frame(), orgba) >= PSNR_MIN
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_train_shards_sum_to_whole(N, ctx):
    """Data-parallel training: the gradients of 2 strided shards add up to the single-GPU gradients (fp32 reassociation
    tolerance), and the shards' records interleave to the whole record list bit for bit."""
    tw, th = 16, 10
    cfg = Cm.scene_cfg()
    r, nrc, m, vol = build(N, ctx, cfg, 32, 32, tw, th)
    r.FrameTrain(); ctx.sync()
    rec_full, cnt_full = r.read_train_probes()
    g_full = nrc.get_params()[1].copy()
    t_full = m.get_tables()[1][:2 * 16384].copy()
    loss_full = nrc.GetStats().mseLoss
    g_sum, t_sum, loss_sum = np.zeros_like(g_full), np.zeros_like(t_full), 0.0
    rec = np.zeros_like(rec_full); cnt = np.zeros_like(cnt_full)
    for k in range(2):
        nrc.set_state(delta=np.zeros(20995, np.float32)); m.set_tables(delta=np.zeros(m.count, np.float32)); nrc.ResetStats()
        r.SetPartitionStrided(0, 1, k, 2); r.FrameTrain(); ctx.sync()
        rk, ck = r.read_train_probes()
        rec[k::2], cnt[k::2] = rk[k::2], ck[k::2]
        g_sum += nrc.get_params()[1]; t_sum += m.get_tables()[1][:2 * 16384]; loss_sum += nrc.GetStats().mseLoss
    np.testing.assert_array_equal(rec.view(np.uint32), rec_full.view(np.uint32))
    np.testing.assert_array_equal(cnt, cnt_full)
    assert _relmax(g_sum, g_full) < 1e-5
    assert _relmax(t_sum, t_full) < 1e-5
    assert abs(loss_sum - loss_full) <= 1e-5 * abs(loss_full) + 1e-12
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_pool_overflow_is_redone(N, ctx):
    """A render vertex pool far below what the view needs overflows: the frame handed out must be bit-identical to the one
    traced with the worst-case pool (the truncated frame is redone before any read-back returns)."""
    W, H = 64, 48
    cfg = Cm.scene_cfg(aspect=W / H)
    r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
    r.SetPoolCapacity(W * H * 32)
    r.FrameRender(); ctx.sync()
    want = r.read_frame().copy()
    cnt_want, qry_want = r.read_render_probes()
    need = int(cnt_want[..., 1].sum())
    assert need > 64, "the view must light more than 64 vertices for this test to bite"
    r.SetPoolCapacity(64)
    r.FrameRender()                      # overflows; nothing is checked yet
    got = r.read_frame().copy()          # read-back settles the check and redoes the frame
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
    cnt, qry = r.read_render_probes()
    np.testing.assert_array_equal(cnt, cnt_want)
    # the pool stays grown: the next frame is right at once
    r.FrameRender(); ctx.sync()
    np.testing.assert_array_equal(r.read_frame().view(np.uint32), want.view(np.uint32))
    # and the nrc_sync path settles it too
    r.SetPoolCapacity(64)
    r.FrameRender(); ctx.sync()
    np.testing.assert_array_equal(np.asarray(r.read_frame()).view(np.uint32), want.view(np.uint32))
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_queued_frames_equal_frames_read_one_by_one(N, ctx):
    """The frame graph runs several frames deep (train trace one frame ahead, double-buffered pools): queueing four
    Render() calls without any host synchronisation must leave the same weights, loss and last frame as four synchronous ones."""
    W, H, tw, th = 64, 48, 16, 16
    cfg = Cm.scene_cfg(aspect=W / H)
    rnd = [(0.37, 0.61, 0.13, 0.89), (0.11, 0.93, 0.5, 0.2), (0.71, 0.29, 0.4, 0.6), (0.05, 0.45, 0.85, 0.25)]
    out = []
    for queued in (False, True):
        r, nrc, m, vol = build(N, ctx, cfg, W, H, tw, th)
        for f in range(4):
            vol.random = np.asarray(rnd[f], np.float32)
            r.Render()
            if not queued:
                ctx.sync()
                r.read_frame()
        ctx.sync()
        out.append((r.read_frame().copy(), nrc.get_params()[0].copy(), m.get_tables()[0].copy(), nrc.GetStats().mseLoss))
        r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()
    assert Cm.psnr(out[1][0], out[0][0]) >= PSNR_MIN
    # gradient accumulation uses float atomics (order varies run to run): weights agree to fp32 reassociation, not bitwise
    assert _relmax(out[1][1], out[0][1]) < 1e-4
    assert _relmax(out[1][2], out[0][2]) < 1e-4
    assert abs(out[0][3] - out[1][3]) <= 1e-3 * abs(out[0][3])


def test_fused_reduce_and_steps_launch_equals_the_three_launches(N, ctx, monkeypatch):
    """NRC_FUSED_STEP=1: Render() sums the train kernel's partial gradients and applies nrc-step + mrhe-step in ONE launch (from the
    second frame on, once the table tail is a fixed point).  Same weights, tables, momentum and frames as the three separate launches,
    up to the summation order inside the train kernel's shared-memory histogram."""
    W, H, tw, th = 64, 48, 16, 16
    cfg = Cm.scene_cfg(aspect=W / H)
    rnd = [(0.37, 0.61, 0.13, 0.89), (0.11, 0.93, 0.5, 0.2), (0.71, 0.29, 0.4, 0.6), (0.05, 0.45, 0.85, 0.25)]
    out = []
    for fused in ("0", "1"):
        monkeypatch.setenv("NRC_FUSED_STEP", fused)
        r, nrc, m, vol = build(N, ctx, cfg, W, H, tw, th)
        n0 = ctx.launch_count()
        for f in range(4):
            vol.random = np.asarray(rnd[f], np.float32)
            r.Render()
        ctx.sync()
        launches = ctx.launch_count() - n0
        p, d, mo = nrc.get_params()
        t, td = m.get_tables()
        assert not d.any() and not td.any()
        out.append((p.copy(), mo.copy(), t.copy(), r.read_frame().copy(), nrc.GetStats().mseLoss, launches))
        r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()
    monkeypatch.delenv("NRC_FUSED_STEP")
    assert out[0][5] - out[1][5] == 2 * 3            # frames 2-4 save two launches each
    for k in (0, 1, 2):
        assert _relmax(out[1][k], out[0][k]) < 1e-5, k
    assert Cm.psnr(out[1][3], out[0][3]) >= 80.0
    assert abs(out[1][4] - out[0][4]) <= 1e-4 * abs(out[0][4])


def test_async_readback_delivers_the_same_frames(N, ctx):
    """Render(k); async copy of frame k; Render(k+1) queued before frame k is awaited: each delivered frame and loss equals
    what the synchronous calls return for the same frame (the next composite must not overwrite a frame under its copy)."""
    import torch
    W, H, tw, th = 64, 48, 16, 16
    cfg = Cm.scene_cfg(aspect=W / H)
    rnd = [(0.37, 0.61, 0.13, 0.89), (0.11, 0.93, 0.5, 0.2), (0.71, 0.29, 0.4, 0.6)]
    r, nrc, m, vol = build(N, ctx, cfg, W, H, tw, th)
    p0, t0 = nrc.get_params()[0].copy(), m.get_tables()[0].copy()
    want, want_loss = [], []
    for f in range(3):
        vol.random = np.asarray(rnd[f], np.float32)
        nrc.ResetStats(); r.Render(); ctx.sync()
        want.append(r.read_frame(N._lib.FRAME_BGRA8).copy()); want_loss.append(nrc.GetStats().mseLoss)
    # same three frames again from the same start, delivered asynchronously
    nrc.set_params(p0); nrc.set_state(delta=np.zeros(20995, np.float32), momentum=np.zeros(20995, np.float32)); m.set_tables(t0)
    bufs = [torch.empty(H * W * 4, dtype=torch.uint8).pin_memory() for _ in range(3)]
    loss = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(3)]
    tickets = []
    for f in range(3):
        vol.random = np.asarray(rnd[f], np.float32)
        nrc.ResetStats(); r.Render()
        nrc.get_stats_async(loss[f].numpy())
        tickets.append(r.read_frame_partition_async(bufs[f].numpy(), N._lib.FRAME_BGRA8))
    for f in range(3):
        r.readback_wait(tickets[f])
        got = bufs[f].numpy().reshape(H, W, 4)
        # 8-bit frames: identical up to the last-bit effects of atomically accumulated gradients on later frames
        assert int(np.abs(got.astype(np.int32) - want[f].astype(np.int32)).max()) <= 1, f
        assert abs(float(loss[f][0]) - want_loss[f]) <= 1e-3 * abs(want_loss[f]) + 1e-12
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


def test_env_importance_half_matches_oracle(N, ctx):
    """Opt-in estimator: with the CDF half of the env gather switched on, the walk's control flow is
    unchanged (same 16 draws per vertex) and the frame matches the oracle run with the same switch."""
    W, H = 64, 48
    cfg = Cm.scene_cfg(use_nn=0, aspect=W / H)
    env_rgba = Cm.envmap()
    eh, ew = env_rgba.shape[:2]
    cdf_x, cdf_y = N.hdr4f_to_cdf(ctx, env_rgba, ew, eh)
    sc = Cm.oracle_scene(cfg)
    sc.set_envmap(env_rgba, cdf_x, cdf_y, cfg["env_direct"], cfg["env_hpm"])
    base, cnt0, _ = sc.render(W, H)
    sc.set_env_cdf_half(True)
    want, cnt1, _ = sc.render(W, H)
    for col in (0, 1, 3):
        np.testing.assert_array_equal(cnt0[..., col], cnt1[..., col])   # same RNG consumption, lit vertices and walk steps
    assert np.isfinite(want).all()
    assert Cm.psnr(want, base) < 50.0                            # ... but a different estimator
    vol = N.VolumeData(ctx, Cm.density())
    vol.useNN, vol.showNonNN, vol.densityFactor, vol.g = 0, 0, cfg["density_factor"], cfg["g"]
    vol.random = cfg["random4"]
    env = N.HdrEnvMap(ctx, ew, eh, env_rgba, cdf_x, cdf_y)
    env.set_strengths(cfg["env_direct"], cfg["env_hpm"])
    dl = N.DirLight(Cm.DIR_ZENITH, Cm.DIR_AZIMUTH, (1, 1, 1), cfg["dir_strength"])
    pl = N.PointLight(cfg["pt_pos"], cfg["pt_color"], cfg["pt_strength"])
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0)
    r = N.NrcHpmRenderer(W, H, 8, 8, cfg["cam"], vol, dl, pl, env, nrc, m)
    r.FrameRender(); ctx.sync()
    assert Cm.psnr(r.read_frame(), base) >= PSNR_MIN             # off by default
    env.set_importance_sampling(True)
    r.FrameRender(); ctx.sync()
    cnt, _ = r.read_render_probes()
    np.testing.assert_array_equal(cnt[..., 0], cnt1[..., 0])
    assert Cm.psnr(r.read_frame(), want) >= PSNR_MIN
    plain = N.HdrEnvMap(ctx, ew, eh, env_rgba)
    with pytest.raises(N.NrcError):
        plain.set_importance_sampling(True)                      # no inverse CDFs: refused
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy(); env.Destroy(); plain.Destroy()


def test_train_env_light_matches_oracle(N, ctx):
    """Opt-in estimator: env-map term in the training targets.  RNG draws, lit vertices, walk steps and the
    records' positions/directions bit for bit; targets to 1e-4; then back to the shipped estimator."""
    tw, th = 20, 12
    cfg = Cm.scene_cfg()
    sc = Cm.oracle_scene(cfg)
    sc.set_train_size(tw, th)
    orec0, ocnt0 = sc.train_trace(tw, th)
    sc.set_train_env_light(True)
    orec, ocnt = sc.train_trace(tw, th)
    r, nrc, m, vol = build(N, ctx, cfg, 32, 32, tw, th)
    r.SetTrainEnvLight(True)
    r.FrameTrain(); ctx.sync()
    rec, cnt = r.read_train_probes()
    np.testing.assert_array_equal(cnt[:, 0], ocnt[:, 0])
    np.testing.assert_array_equal(cnt[:, 1], ocnt[:, 1])
    np.testing.assert_array_equal(cnt[:, 2], ocnt[:, 3])
    np.testing.assert_array_equal(rec[:, :6].view(np.uint32), orec[:, :6].view(np.uint32))
    assert _relmax(rec[:, 6:], orec[:, 6:]) < 1e-4
    assert float(rec[:, 6:].sum()) > float(orec0[:, 6:].sum())          # the env term adds light
    # two queued frames with the flag on (double-buffered pools carry the random numbers), then off again
    r.FrameTrain(); r.FrameTrain(); ctx.sync()
    r.SetTrainEnvLight(False)
    nrc.set_state(delta=np.zeros(20995, np.float32)); nrc.ResetStats()
    r.FrameTrain(); ctx.sync()
    rec, cnt = r.read_train_probes()
    np.testing.assert_array_equal(cnt[:, 0], ocnt0[:, 0])
    assert _relmax(rec[:, 6:], orec0[:, 6:]) < 1e-4
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()
