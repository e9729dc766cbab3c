"""-m gpu: the tcgen05 (bf16 operands, fp32 accumulate in TMEM) MLP path against the fp32 oracle.
Stated bound for this path (north_star: "a stated looser bound for any bf16 tensor-core path"): max abs error of the
network output <= TC_REL of the largest output magnitude of the batch; hash indices stay bit-exact (same code)."""
import numpy as np
import pytest

import common as Cm

pytestmark = pytest.mark.gpu

TC_REL = 3e-2


@pytest.fixture(scope="module")
def N():
    import neuralradiancecaching_b200 as N
    return N


@pytest.fixture()
def ctx(N):
    c = N.Context(0)
    c.set_precision(N._lib.MLP_BF16_TC)
    yield c
    c.destroy()


def _rel(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) / max(float(np.max(np.abs(b))), 1e-30))


def _pair(N, ctx, wscale=0.01):
    params = Cm.S.mlp_init(1)
    params[:20672] *= wscale / 0.01
    params[20672:] = Cm.S.init_params(7, 323, 0.01)
    tables = Cm.S.table_init()
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=params, tables=tables)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    nrc.set_params(params)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0)
    m.set_tables(tables)
    return sc, nrc, m


@pytest.mark.parametrize("wscale", [0.01, 0.1])
def test_infer_tc_matches_oracle(N, ctx, wscale):
    sc, nrc, m = _pair(N, ctx, wscale)
    for n in (1, 127, 128, 129, 5000, 1 << 16):
        pos, d = Cm.rand_queries(n, seed=300 + n)
        got = N.infer(nrc, m, pos, d)
        want = sc.infer(pos, d)
        assert np.isfinite(got).all()
        assert _rel(got, want) < TC_REL, (n, _rel(got, want))
    nrc.Destroy(); m.Destroy()


def test_infer_tc_after_step_uses_new_weights(N, ctx):
    sc, nrc, m = _pair(N, ctx)
    pos, d = Cm.rand_queries(2048, seed=9)
    N.infer(nrc, m, pos, d)                      # builds the bf16 operand image
    target = np.abs(pos) / 40.0
    sc.set_loss_scale(1.0 / 2048, 1.0)
    sc.train_accumulate(pos, d, target)
    N.train_accumulate(nrc, m, pos, d, target, 1.0 / 2048)
    nrc.set_config(0.05, 0.0, 0.5); sc.set_mlp(None, 0.05, 0.0, 0.5)
    N.optimizer_step(nrc, m); sc.nrc_step(); sc.mrhe_step()
    got, want = N.infer(nrc, m, pos, d), sc.infer(pos, d)
    assert _rel(got, want) < TC_REL
    nrc.Destroy(); m.Destroy()


def test_frame_with_tc_inference(N, ctx):
    from test_gpu_frame import build
    W, H = 96, 64
    cfg = Cm.scene_cfg(aspect=W / H)
    sc = Cm.oracle_scene(cfg)
    orgba, _, oqry = sc.render(W, H)
    r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
    r.FrameRender(); ctx.sync()
    frame = r.read_frame()
    assert r.stats()["infer_queries"] == int(oqry[..., 7].sum())     # control flow is unaffected by the MLP precision
    assert Cm.psnr(frame, orgba) >= 40.0                              # stated PSNR for the bf16 path (fp32 path: 60 dB)
    r.Destroy(); nrc.Destroy(); m.Destroy(); vol.Destroy()


# Stated bounds for the tensor-core TRAINING path (per tensor, max abs error relative to the tensor's largest magnitude).  Round 1 ran
# the whole kernel in bf16 and measured 7.6 % on the MLP gradients and 30 % on the hash-table gradient: the ReLU gates of the backward
# pass were read off bf16-rounded forward activations (sign flips of pre-activations inside the rounding noise; OneBlob bins that are
# fp32 denormals round to +0 in bf16 and closed the layer-0 gate for ~20 % of all directions), and the reference's unbounded corner
# weights let a few such samples dominate a table entry.  Now the forward pass runs in split precision (bf16x3, six
# partial products per K step: fp32-accurate activations and gates), the layer-0 gate is taken from the exact fp32 inputs, and only the
# backward GEMMs stay bf16: measured 0.2-0.5 % on the MLP gradients, 0.4-0.7 % on the table gradient (scripts/tc_err.py,
# scripts/tc_split_emul.py).  The fp32 path keeps the 1e-3 bar for all of these.
TC_GRAD_REL = 3e-2
TC_TABLE_GRAD_REL = 2e-2
TC_COSINE_MIN = 0.999          # direction-sensitive: per layer / per level, a wrong sign or scale of one block cannot hide in a max-norm


def _cosine(a, b):
    a, b = a.astype(np.float64).ravel(), b.astype(np.float64).ravel()
    na, nb = np.linalg.norm(a), np.linalg.norm(b)
    return 1.0 if na == 0 and nb == 0 else float(a @ b / max(na * nb, 1e-300))


# the forms of csrc/mlp_tc.cu::k_train_tc: two / four threads per sample row, and the warp-specialised one (encoder warps a tile ahead); the
# launcher picks by batch size
@pytest.mark.parametrize("form", ["256", "512", "ws"])
@pytest.mark.parametrize("n", [128, 3000, 1 << 15])
def test_train_tc_gradients_match_oracle(N, ctx, n, form, monkeypatch):
    monkeypatch.setenv("NRC_TRAIN_TC_FORM", form)
    sc, nrc, m = _pair(N, ctx, 0.1)
    pos, d = Cm.rand_queries(n, seed=21 + n)
    g = np.random.Generator(np.random.PCG64(4))
    target = (g.random((n, 3)) * 3.0).astype(np.float32)
    if n > 100:
        target[5] = [2000.0, 0.5, 0.1]      # the 1024 clamp
        pos[7] = [np.nan, 0.0, 0.0]         # the NaN guard
    scale = 1.0 / n
    sc.set_loss_scale(scale, 1.0)
    opred, dropped = sc.train_accumulate(pos, d, target)
    gpred = N.train_accumulate(nrc, m, pos, d, target, scale, want_pred=True)
    assert np.isfinite(gpred).all()
    assert _rel(gpred, opred) < TC_REL
    _, od, _ = sc.get_mlp()
    _, gd, _ = nrc.get_params()
    assert np.isfinite(gd).all()
    for l in range(6):
        a, b = (l * 4096, l * 4096 + (4096 if l < 5 else 192))
        assert _rel(gd[a:b], od[a:b]) < TC_GRAD_REL, ("dW%d" % l, _rel(gd[a:b], od[a:b]))
        assert _cosine(gd[a:b], od[a:b]) >= TC_COSINE_MIN, ("dW%d cosine" % l, _cosine(gd[a:b], od[a:b]))
        a, b = (20672 + 64 * l, 20672 + 64 * l + (64 if l < 5 else 3))
        assert _rel(gd[a:b], od[a:b]) < TC_GRAD_REL, ("db%d" % l, _rel(gd[a:b], od[a:b]))
        assert _cosine(gd[a:b], od[a:b]) >= TC_COSINE_MIN, ("db%d cosine" % l)
    _, otd = sc.get_mrhe()
    _, gtd = m.get_tables()
    assert np.count_nonzero(gtd[2 * 16384:]) == 0
    assert _rel(gtd[:2 * 16384], otd[:2 * 16384]) < TC_TABLE_GRAD_REL
    assert _cosine(gtd[:2 * 16384], otd[:2 * 16384]) >= TC_COSINE_MIN
    # both features of an entry receive the same value (A9), as in the oracle
    np.testing.assert_array_equal(gtd[0:2 * 16384:2], gtd[1:2 * 16384:2])
    assert abs(nrc.GetStats().mseLoss - sc.mse_loss()) <= TC_REL * abs(sc.mse_loss())
    nrc.Destroy(); m.Destroy()


def test_train_tc_forms_agree_over_many_tiles_per_cta(N, ctx, monkeypatch):
    """More than two tiles per CTA (148 SMs x 128 samples x 3 and a ragged end): the warp-specialised form re-uses its two input sets
    (the EMPTY / FULL hand-over), the dW accumulators stay in TMEM over all of a CTA's tiles.  Same tile-to-CTA mapping in all forms, so
    predictions and MLP gradients must agree to the last bit with the 256-thread form (itself checked against the oracle above); the
    hash-table gradient differs by the order of the shared-memory atomic adds only."""
    n = 148 * 128 * 3 + 1234
    pos, d = Cm.rand_queries(n, seed=5)
    target = (np.random.Generator(np.random.PCG64(6)).random((n, 3)) * 2.0).astype(np.float32)
    pos[123] = [np.nan, 1.0, 2.0]
    got = {}
    for form in ("256", "512", "ws", "ws"):
        monkeypatch.setenv("NRC_TRAIN_TC_FORM", form)
        _, nrc, m = _pair(N, ctx, 0.1)
        pred = N.train_accumulate(nrc, m, pos, d, target, 1.0 / n, want_pred=True)
        got.setdefault(form, []).append((pred.copy(), nrc.get_params()[1].copy(), m.get_tables()[1][:2 * 16384].copy(), nrc.GetStats().mseLoss))
        nrc.Destroy(); m.Destroy()
    ref = got["256"][0]
    assert np.isfinite(ref[1]).all() and np.abs(ref[1]).max() > 0
    for form in ("512", "ws"):
        for run in got[form]:
            np.testing.assert_array_equal(run[0].view(np.uint32), ref[0].view(np.uint32))
            np.testing.assert_array_equal(run[1].view(np.uint32), ref[1].view(np.uint32))
            assert _rel(run[2], ref[2]) < 1e-5, (form, _rel(run[2], ref[2]))
            assert run[3] == ref[3]


@pytest.mark.parametrize("form", ["256", "512", "ws"])
def test_tc_training_converges_like_fp32(N, ctx, form, monkeypatch):
    """40 optimiser steps on a fixed batch: the loss under the tensor-core path tracks the fp32 oracle's step by step."""
    monkeypatch.setenv("NRC_TRAIN_TC_FORM", form)
    sc, nrc, m = _pair(N, ctx, 0.1)
    n = 4096
    pos, d = Cm.rand_queries(n, seed=77)
    target = (np.abs(pos) / 40.0).astype(np.float32)
    sc.set_loss_scale(1.0 / n, 1.0)
    nrc.set_config(0.02, 0.0, 0.5); sc.set_mlp(None, 0.02, 0.0, 0.5)
    lo, lg = [], []
    for it in range(40):
        sc.reset_stats(); nrc.ResetStats()
        sc.train_accumulate(pos, d, target)
        N.train_accumulate(nrc, m, pos, d, target, 1.0 / n)
        lo.append(sc.mse_loss()); lg.append(nrc.GetStats().mseLoss)
        sc.nrc_step(); sc.mrhe_step(); N.optimizer_step(nrc, m)
    assert lo[-1] < 0.7 * lo[0]                      # it trains
    assert max(abs(a - b) / b for a, b in zip(lg, lo)) <= 0.03      # and the tensor-core path follows the fp32 loss curve within 3 % at every step
    nrc.Destroy(); m.Destroy()


# ---- split inference: k_encode_tiles -> bf16 operand-image tiles -> k_mlp_stream (pipelined, A operand from TMEM) ----------
@pytest.mark.parametrize("wscale", [0.01, 0.1])
def test_infer_split_matches_oracle_and_fused(N, ctx, wscale):
    sc, nrc, m = _pair(N, ctx, wscale)
    # 1<<18 queries = 2048 tiles: ~14 tiles per CTA, so every shared-memory stage and every TMEM slot is reused
    for n in (1, 127, 128, 129, 5000, 148 * 128 * 5 + 77, 1 << 18):
        pos, d = Cm.rand_queries(n, seed=500 + n)
        ctx.set_infer_mode(False)
        fused = N.infer(nrc, m, pos, d)
        ctx.set_infer_mode(True)
        got = N.infer(nrc, m, pos, d)
        assert np.isfinite(got).all()
        # same operands, same fp32 accumulation: the two kernels may differ only where bf16 rounding of an activation flips
        assert _rel(got, fused) < 1e-6, (n, _rel(got, fused))
        if n <= (1 << 16):
            want = sc.infer(pos, d)
            assert _rel(got, want) < TC_REL, (n, _rel(got, want))
    ctx.set_infer_mode(False)
    nrc.Destroy(); m.Destroy()


@pytest.mark.parametrize("slots", [1, 2, 3, 4, 5])
def test_mlp_stream_slots_agree(N, ctx, slots, monkeypatch):
    import torch
    sc, nrc, m = _pair(N, ctx, 0.1)
    n = 148 * 128 * 11 + 5
    pos, d = Cm.rand_queries(n, seed=77)
    dpos, dd = torch.from_numpy(pos).cuda(), torch.from_numpy(d).cuda()
    ctx.set_infer_mode(False)
    fused = N.infer(nrc, m, dpos, dd)
    tiles = N.encode_tiles(m, dpos, dd)
    assert tiles.numel() == N.feature_tile_bytes(n) == ((n + 127) // 128) * 16384
    monkeypatch.setenv("NRC_MLP_SLOTS", str(slots))
    got = N.mlp_forward_tiles(nrc, tiles, n, like=dpos)
    ctx.sync()
    assert _rel(got.cpu().numpy(), fused.cpu().numpy()) < 1e-6
    nrc.Destroy(); m.Destroy()


def test_encode_tiles_layout(N, ctx):
    """chunk (kc, row) of a tile holds network inputs 8kc..8kc+7 of query row, bf16 of the exact fp32 features."""
    import torch
    sc, nrc, m = _pair(N, ctx)
    n = 300
    pos, d = Cm.rand_queries(n, seed=3)
    tiles = N.encode_tiles(m, torch.from_numpy(pos).cuda(), torch.from_numpy(d).cuda())
    ctx.sync()
    raw = tiles.cpu().numpy().view(np.uint16).reshape(-1, 8, 128, 8)          # [tile][kc][row][8]
    feat = np.concatenate([N.encode(m, pos, want_idx=False)[0], N.encode_dir(ctx, d)], axis=1)    # [n][64] fp32
    u = feat.view(np.uint32).astype(np.uint64)
    want = (((u + 0x7FFF + ((u >> 16) & 1)) >> 16) & 0xFFFF).astype(np.uint16)     # round-to-nearest-even bf16 (finite inputs)
    got = raw.transpose(0, 2, 1, 3).reshape(-1, 64)
    assert np.array_equal(got[:n], want)
    assert not got[n:].any()
    nrc.Destroy(); m.Destroy()


def test_frame_split_equals_fused(N, ctx):
    from test_gpu_frame import build
    W, H = 96, 64
    cfg = Cm.scene_cfg(aspect=W / H)
    frames = []
    for split in (False, True):
        ctx.set_infer_mode(split)
        r, nrc, m, vol = build(N, ctx, cfg, W, H, 8, 8)
        r.FrameRender(); ctx.sync()          # no training step: both runs infer with identical weights
        frames.append(r.read_frame())
        r.Destroy()
    ctx.set_infer_mode(False)
    assert Cm.psnr(frames[1], frames[0]) >= 80.0
