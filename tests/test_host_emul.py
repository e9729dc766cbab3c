"""The device walk / encode code (csrc/strict.cuh, csrc/walk.cuh) compiled for the HOST by tests/host_emul and compared
bit for bit with the oracle.  This is how bit-exact control flow is debugged on the GPU-less build machine; the
-m gpu tests repeat the comparison through the real kernels."""
import numpy as np

import common as Cm
from host_emul import emul


def test_render_walk_bit_exact():
    cfg = Cm.scene_cfg()
    sc = Cm.oracle_scene(cfg)
    shape = emul.set_volume(Cm.density())
    W, H = 64, 40
    _, cnt, qry = sc.render(W, H)
    ecnt, eqry = emul.render_walk(cfg, W, H, shape)
    np.testing.assert_array_equal(cnt[..., 0], ecnt[..., 0])   # RNG draws
    np.testing.assert_array_equal(cnt[..., 1], ecnt[..., 1])   # lit vertices
    np.testing.assert_array_equal(cnt[..., 3], ecnt[..., 2])   # walk steps
    np.testing.assert_array_equal(qry[..., 7], eqry[..., 6])
    np.testing.assert_array_equal(qry[..., :6].view(np.uint32), eqry[..., :6].view(np.uint32))
    assert qry[..., 7].mean() > 0.2   # the scene does exercise the cache


def test_render_walk_in_two_launches_is_the_same_walk():
    """trace.cu stops the render walk after NRC_RWALK_SPLIT steps and finishes the surviving paths in a second launch from a 20-word
    record: the record carries everything the remaining steps read (every split point gives the single walk's draws, vertices,
    steps and queries), for the cache on and off."""
    shape = emul.set_volume(Cm.density())
    W, H = 64, 40
    for cfg in (Cm.scene_cfg(), Cm.scene_cfg(use_nn=0)):
        cnt, qry = emul.render_walk(cfg, W, H, shape)
        handed = []
        for k in (1, 4, 10, 31):
            scnt, sqry, n = emul.render_walk_split(cfg, W, H, shape, k)
            np.testing.assert_array_equal(scnt, cnt)
            np.testing.assert_array_equal(sqry.view(np.uint32), qry.view(np.uint32))
            handed.append(n)
        assert handed[0] >= handed[1] >= handed[2] >= handed[3] > 0          # fewer paths survive a later split
        assert handed[0] == int((cnt[..., 2] > 1).sum())                      # exactly the paths with a second step


def test_walk_cost_counters_describe_the_same_walk():
    """the per-step SDF counts behind scripts/walk_divergence.py come from a step-by-step copy of the walk loop: its steps and lit
    vertices must be those of render_walk_path, and the per-step counts must add up to the walk's total"""
    cfg = Cm.scene_cfg()
    shape = emul.set_volume(Cm.density())
    W, H = 48, 30
    cnt, _ = emul.render_walk(cfg, W, H, shape)
    cost, per = emul.render_walk_cost(cfg, W, H, shape)
    np.testing.assert_array_equal(cost[..., 0], cnt[..., 2])                 # walk steps
    np.testing.assert_array_equal(cost[..., 3], cnt[..., 1])                 # lit vertices
    np.testing.assert_array_equal(per.sum(axis=2), cost[..., 1])             # SDF evaluations, step by step
    assert (per[(cost[..., 0] == 32) & (cnt[..., 3] == 0)] > 0).all()         # a walk that runs out of steps marched in every one


def test_render_walk_frame0_and_no_nn():
    # frame 0: random = 0 -> every draw is 0; and the non-NN path (useNN = 0)
    for cfg in (Cm.scene_cfg(random4=(0, 0, 0, 0)), Cm.scene_cfg(use_nn=0)):
        sc = Cm.oracle_scene(cfg)
        shape = emul.set_volume(Cm.density())
        W, H = 32, 20
        _, cnt, qry = sc.render(W, H)
        ecnt, eqry = emul.render_walk(cfg, W, H, shape)
        np.testing.assert_array_equal(cnt[..., [0, 1, 3]], ecnt[..., [0, 1, 2]])
        np.testing.assert_array_equal(qry[..., :6].view(np.uint32), eqry[..., :6].view(np.uint32))


def test_train_walk_bit_exact():
    cfg = Cm.scene_cfg()
    sc = Cm.oracle_scene(cfg)
    shape = emul.set_volume(Cm.density())
    tw, th = 12, 10
    sc.set_train_size(tw, th)
    rec, cnt = sc.train_trace(tw, th)
    erec, ecnt = emul.train_walk(cfg, tw, th, shape)
    np.testing.assert_array_equal(cnt[:, [0, 1, 3]], ecnt)
    np.testing.assert_array_equal(rec[:, :6].view(np.uint32), erec.view(np.uint32))


def test_train_walk_with_env_light_bit_exact():
    # opt-in: env-map term in the training targets -> 16 more draws per lit vertex, in the render walk's order
    cfg = Cm.scene_cfg()
    sc = Cm.oracle_scene(cfg)
    shape = emul.set_volume(Cm.density())
    tw, th = 12, 10
    sc.set_train_size(tw, th)
    rec0, cnt0 = sc.train_trace(tw, th)
    sc.set_train_env_light(True)
    rec, cnt = sc.train_trace(tw, th)
    assert (cnt[:, 0] > cnt0[:, 0]).any()                       # more RNG draws than the shipped estimator
    erec, ecnt = emul.train_walk(cfg, tw, th, shape, env_light=True)
    np.testing.assert_array_equal(cnt[:, [0, 1, 3]], ecnt)
    np.testing.assert_array_equal(rec[:, :6].view(np.uint32), erec.view(np.uint32))
    assert np.isfinite(rec).all() and (rec[:, 6:9] >= rec0[:, 6:9] - 1e-6).mean() > 0.5


def test_encode_bit_exact_both_resolution_tables():
    pos, d = Cm.rand_queries(2048, seed=3)
    pos[:8] = [[0, 0, 0], [31.25, 21.25, 38.25], [-31.25, -21.25, -38.25], [-31.4, 0, 0], [0, -21.4, 38.3], [31.2499, 21.2499, 38.2499],
               [1e-30, -1e-30, 0], [10, 10, 10]]
    tables = Cm.S.table_init()
    sc = Cm.oracle_scene(Cm.scene_cfg())
    for res in (Cm.RES_INTENDED, Cm.RES_SHIPPED):
        sc.set_mrhe(tables, res)
        f, i, _ = sc.encode_pos(pos)
        ef, ei = emul.encode(pos, tables, 16, 16384, res)
        np.testing.assert_array_equal(i, ei)
        np.testing.assert_array_equal(f.view(np.uint32), ef.view(np.uint32))
    np.testing.assert_array_equal(sc.encode_dir(d).view(np.uint32), emul.oneblob(d).view(np.uint32))
