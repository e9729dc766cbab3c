"""CPU: the opt-in "fixed" preset of the oracle -- properties the corrected formulas must have and the shipped
ones do not.  The parity preset itself is pinned by test_oracle_kat.py; this file pins the switch."""
import numpy as np

import common as Cm


def _params():
    p = Cm.S.mlp_init(1)
    p[:20672] *= 12.0          # N(0, 0.12): about 1/sqrt(64), so that a table entry's influence survives six layers
    p[20672:] = 0.05           # keeps the ReLUs open
    return p


def _scene(fixed, T=16384):
    tables = Cm.S.table_init(2, 16, T, 2)
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=_params(), tables=tables, T=T)
    sc.set_mrhe_fixed(fixed)
    return sc, tables


def test_fixed_encode_interpolates_the_grid():
    """at a grid vertex of level l the fixed encoder returns that vertex's table entry (factors are 0) -- the shipped one does not"""
    sc, tables = _scene(True)
    res = Cm.RES_INTENDED
    sky = np.array([62.5, 42.5, 76.5], np.float32)
    level = 3
    r = int(res[level])
    v = np.array([5, 7, 9], np.float32)                       # grid vertex
    pos = ((v / r - 0.5) * sky).astype(np.float32)[None, :]     # normPos * res == v up to rounding: nudge inside the cell
    pos += 1e-4
    feat, idx, lerp = sc.encode_pos(pos)
    assert np.all(lerp[0, level] >= 0) and np.all(lerp[0, level] < 0.05)
    t = tables.reshape(16, -1, 2)
    h = (5 * 1 + 7 * 19349663 + 9 * 83492791) % (2 ** 32) % 16384
    assert idx[0, level * 8 + 0] == h
    np.testing.assert_allclose(feat[0, 2 * level:2 * level + 2], t[level, h], rtol=0, atol=0.05 * np.abs(t[level]).max() * 3)
    # every fixed factor is a fraction in [0, 1); the shipped factors (world minus grid) are not
    pos2, _ = Cm.rand_queries(512, seed=5)
    _, _, lf = sc.encode_pos(pos2)
    assert lf.min() >= 0.0 and lf.max() < 1.0
    sc0, _ = _scene(False)
    _, _, l0 = sc0.encode_pos(pos2)
    assert l0.min() < 0.0 or l0.max() >= 1.0


def test_fixed_backward_scatters_per_level_per_feature_with_partition_of_unity_weights():
    """one sample: every level's own table receives 8 corner gradients whose weights are the products of the (grid-space) factors
    and sum to one, and feature 1 has its own error (the MLP part of Backprop keeps the reference's gating)"""
    sc, tables = _scene(True)
    pos, d = Cm.rand_queries(1, seed=11)
    target = (np.abs(pos) / 40.0).astype(np.float32)
    sc.set_loss_scale(1.0, 1.0)
    sc.train_accumulate(pos, d, target)
    _, delta = sc.get_mrhe()
    _, idx, lf = sc.encode_pos(pos)
    dl = delta.reshape(16, 16384, 2)
    ratios = []
    for level in range(16):
        touched = np.flatnonzero(dl[level].any(axis=1))
        corners = idx[0, level * 8:level * 8 + 8]
        assert set(touched) <= set(int(c) for c in corners)        # nothing outside this level's 8 corners (A10 fixed)
        if len(set(int(c) for c in corners)) < 8:
            continue                                                # hash collision between corners: sums merge
        x, y, z = lf[0, level]
        for f in (0, 1):
            tot = dl[level, corners, f].sum()
            if abs(tot) < 1e-12:
                continue
            for c in range(8):
                w = (x if c & 4 else 1 - x) * (y if c & 2 else 1 - y) * (z if c & 1 else 1 - z)
                assert abs(dl[level, corners[c], f] / tot - w) < 1e-4, (level, f, c)
        if abs(dl[level, corners, 0].sum()) > 1e-12:
            ratios.append(dl[level, corners, 1].sum() / dl[level, corners, 0].sum())
    assert len(ratios) >= 8 and np.std(ratios) > 1e-3              # feature 1 is not a copy of feature 0 (A9 fixed)
    # the shipped preset on the same sample: everything lands in level 0's slice, both features identical
    sc0, _ = _scene(False)
    sc0.set_loss_scale(1.0, 1.0)
    sc0.train_accumulate(pos, d, target)
    _, d0 = sc0.get_mrhe()
    assert not d0[2 * 16384:].any()
    d0 = d0[:2 * 16384].reshape(-1, 2)
    np.testing.assert_array_equal(d0[:, 0], d0[:, 1])


def test_adam_first_step_moves_by_lr_along_the_gradient_sign():
    sc, tables = _scene(True)
    sc.set_adam(True, beta2=0.999, eps=1e-8, mrhe_beta1=0.9)
    p0, _, _ = sc.get_mlp()
    g = np.random.Generator(np.random.PCG64(3))
    dp = g.standard_normal(p0.size); dp = (np.sign(dp) * (0.005 + np.abs(dp) * 0.01)).astype(np.float32)   # |g| >> eps
    dp[:5] = 0.0
    sc.set_mlp_state(delta=dp)
    dt = np.zeros_like(tables); x = g.standard_normal(1000); dt[:1000] = (np.sign(x) * (0.005 + np.abs(x) * 0.01)).astype(np.float32)
    sc.set_mrhe_delta(dt)
    sc.nrc_step(); sc.mrhe_step()
    p1, d1, _ = sc.get_mlp()
    t1, td1 = sc.get_mrhe()
    assert not d1.any() and not td1.any()
    # bias-corrected first step: mhat / sqrt(vhat) = g / |g| -> |dw| = lr (0.001 for the MLP, 0.01 for the tables)
    np.testing.assert_allclose(p1[5:] - p0[5:], 0.001 * np.sign(dp[5:]), rtol=2e-3, atol=1e-9)
    np.testing.assert_array_equal(p1[:5], p0[:5])
    np.testing.assert_allclose(t1[:1000] - tables[:1000], 0.01 * np.sign(dt[:1000]), rtol=2e-3, atol=1e-9)
    np.testing.assert_array_equal(t1[1000:], tables[1000:])
