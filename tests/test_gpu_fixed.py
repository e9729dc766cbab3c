"""-m gpu: the opt-in "fixed" preset through the C-ABI against the oracle's switches: corrected MRHE
interpolation / gradient scatter (nrc_mrhe_set_fixed_mode) and Adam in both step kernels (nrc_*_set_adam)."""
import numpy as np
import pytest

import common as Cm

pytestmark = pytest.mark.gpu

REL = 1e-3
TC_REL = 3e-2


@pytest.fixture(scope="module")
def N():
    import neuralradiancecaching_b200 as N
    return N


@pytest.fixture()
def ctx(N):
    c = N.Context(0)
    yield c
    c.destroy()


def _rel(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) / max(float(np.max(np.abs(b))), 1e-30))


def _pair(N, ctx, wscale=1.0):
    params = Cm.S.mlp_init(1)
    params[:20672] *= wscale
    params[20672:] = Cm.S.init_params(7, 323, 0.01)
    tables = Cm.S.table_init(2, 16, 16384, 2)
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=params, tables=tables)
    sc.set_mrhe_fixed(True)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    nrc.set_params(params)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0)
    m.set_tables(tables)
    m.set_fixed_mode(True)
    return sc, nrc, m, params, tables


def test_fixed_encode_is_bit_exact(N, ctx):
    sc, nrc, m, _, _ = _pair(N, ctx)
    pos, _ = Cm.rand_queries(5000, seed=31)
    pos[3] = [31.3, 21.3, 38.3]          # marginally outside the box
    feat, idx = N.encode(m, pos)
    of, oi, _ = sc.encode_pos(pos)
    np.testing.assert_array_equal(idx, oi)
    np.testing.assert_array_equal(feat.view(np.uint32), of.view(np.uint32))
    # and it is a different encoder from the shipped one
    m.set_fixed_mode(False)
    f0, i0 = N.encode(m, pos)
    np.testing.assert_array_equal(i0, oi)                 # same hash indices
    assert not np.array_equal(f0, feat)
    nrc.Destroy(); m.Destroy()


@pytest.mark.parametrize("precision", ["fp32", "bf16", "bf16_split"])
def test_fixed_infer_matches_oracle(N, ctx, precision):
    sc, nrc, m, _, _ = _pair(N, ctx, 12.0)
    if precision != "fp32":
        ctx.set_precision(N._lib.MLP_BF16_TC)
        ctx.set_infer_mode(precision == "bf16_split")
    for n in (1, 129, 5000):
        pos, d = Cm.rand_queries(n, seed=40 + n)
        got, want = N.infer(nrc, m, pos, d), sc.infer(pos, d)
        assert _rel(got, want) < (REL if precision == "fp32" else TC_REL), (n, _rel(got, want))
    nrc.Destroy(); m.Destroy()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_fixed_train_accumulate_matches_oracle(N, ctx, precision):
    """bf16 contexts train the fixed preset on the fp32 kernel (launch_train_tc routes there), so both meet the fp32 bound"""
    sc, nrc, m, _, _ = _pair(N, ctx, 12.0)
    if precision == "bf16":
        ctx.set_precision(N._lib.MLP_BF16_TC)
    n = 3000
    pos, d = Cm.rand_queries(n, seed=51)
    g = np.random.Generator(np.random.PCG64(6))
    target = (g.random((n, 3)) * 3.0).astype(np.float32)
    sc.set_loss_scale(1.0 / n, 1.0)
    opred, _ = sc.train_accumulate(pos, d, target)
    gpred = N.train_accumulate(nrc, m, pos, d, target, 1.0 / n, want_pred=True)
    assert _rel(gpred, opred) < REL
    _, od, _ = sc.get_mlp()
    _, gd, _ = nrc.get_params()
    assert _rel(gd, od) < REL
    _, otd = sc.get_mrhe()
    _, gtd = m.get_tables()
    assert np.count_nonzero(otd[2 * 16384:]) > 0                     # every level trains (A10 fixed)
    for level in range(16):
        a, b = level * 2 * 16384, (level + 1) * 2 * 16384
        assert _rel(gtd[a:b], otd[a:b]) < REL, level
    t = gtd.reshape(-1, 2)
    assert not np.array_equal(t[:, 0], t[:, 1])                      # per-feature errors (A9 fixed)
    assert int(m.grad_ptr()[1]) == 16 * 16384 * 2                    # the data-parallel exchange covers the whole table
    nrc.Destroy(); m.Destroy()


def test_adam_steps_are_bit_exact(N, ctx):
    sc, nrc, m, params, tables = _pair(N, ctx)
    sc.set_adam(True, beta2=0.99, eps=1e-7, mrhe_beta1=0.8)
    nrc.set_adam(True, beta2=0.99, eps=1e-7)
    m.set_adam(True, beta1=0.8, beta2=0.99, eps=1e-7)
    g = np.random.Generator(np.random.PCG64(8))
    for step in range(4):
        dp = (g.standard_normal(params.size) * 0.02).astype(np.float32)
        dt = (g.standard_normal(tables.size) * 0.02).astype(np.float32)
        if step == 1:
            dp[3], dp[20700] = np.nan, np.inf              # scrubbing rules still apply
            dt[5], dt[70000] = -np.inf, 5000.0
        sc.set_mlp_state(delta=dp); sc.set_mrhe_delta(dt)
        nrc.set_state(delta=dp); m.set_tables(delta=dt)
        N.optimizer_step(nrc, m)
        sc.nrc_step(); sc.mrhe_step()
        op, od, om = sc.get_mlp()
        gp, gd, gm = nrc.get_params()
        np.testing.assert_array_equal(gp.view(np.uint32), op.view(np.uint32), err_msg="params, step %d" % step)
        np.testing.assert_array_equal(gm.view(np.uint32), om.view(np.uint32))
        assert not gd.any() and not od.any()
        ot, otd = sc.get_mrhe()
        gt, gtd = m.get_tables()
        np.testing.assert_array_equal(gt.view(np.uint32), ot.view(np.uint32), err_msg="tables, step %d" % step)
        assert not gtd.any()
    # switching Adam off returns to the reference's step, bit for bit
    nrc.set_adam(False); m.set_adam(False); sc.set_adam(False)
    dp = (g.standard_normal(params.size) * 0.02).astype(np.float32)
    dt = (g.standard_normal(tables.size) * 0.02).astype(np.float32)
    sc.set_mlp_state(delta=dp); sc.set_mrhe_delta(dt)
    nrc.set_state(delta=dp); m.set_tables(delta=dt)
    N.optimizer_step(nrc, m); sc.nrc_step(); sc.mrhe_step()
    np.testing.assert_array_equal(nrc.get_params()[0].view(np.uint32), sc.get_mlp()[0].view(np.uint32))
    np.testing.assert_array_equal(m.get_tables()[0].view(np.uint32), sc.get_mrhe()[0].view(np.uint32))
    nrc.Destroy(); m.Destroy()


def test_fixed_preset_fits_better_than_the_shipped_one(N, ctx):
    """the point of the preset: on the same data and step count the corrected encoder + Adam reaches a lower loss"""
    import bench
    n, steps = 1 << 14, 60
    g = np.random.Generator(np.random.PCG64(9))
    pos = ((g.random((n, 3)) - 0.5) * np.array([62.5, 42.5, 76.5])).astype(np.float32)
    d = g.standard_normal((n, 3)); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    target = bench.synthetic_field(pos, d)
    losses = {}
    for fixed in (False, True):
        nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(Cm.S.mlp_init(1))
        m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(Cm.S.table_init(2))
        if fixed:
            m.set_fixed_mode(True); nrc.set_adam(True); m.set_adam(True)
        for _ in range(steps):
            nrc.ResetStats()
            N.train_accumulate(nrc, m, pos, d, target, 1.0 / n)
            N.optimizer_step(nrc, m)
        losses[fixed] = nrc.GetStats().mseLoss
        assert np.isfinite(losses[fixed])
        nrc.Destroy(); m.Destroy()
    assert losses[True] < losses[False], losses
