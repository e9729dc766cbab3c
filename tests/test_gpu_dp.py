"""-m gpu: the fused peer-memory exchange + optimiser steps (csrc/dp.cu) against the oracle's steps on the rank-ordered sum.

The multi-process path needs several GPUs (bench.py --exchange fused under torchrun); here the "ranks" are replicas inside
one process on cuda:0, which runs the same kernel over the same pointer table, only without the cross-GPU flag waits."""
import numpy as np
import pytest

import common as Cm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def N():
    import neuralradiancecaching_b200 as N
    return N


@pytest.fixture(scope="module")
def ctx(N):
    c = N.Context(0)
    yield c
    c.destroy()


def _sum_rank_order(parts):
    acc = np.zeros_like(parts[0])
    for p in parts:                      # fp32 adds in rank order, starting from +0 like the kernel
        acc = (acc + p).astype(np.float32)
    return acc


@pytest.mark.parametrize("world,wd,shard", [(1, 0.0, "0"), (3, 0.0, "0"), (4, 0.01, "0"), (3, 0.0, "1"), (4, 0.01, "1")])
def test_fused_reduce_step_matches_oracle_bit_exact(N, ctx, world, wd, shard, monkeypatch):
    """shard = "1": the sharded form (reduce-scatter + sharded step + all-gather in one kernel; what tables beyond 2^20 head floats
    use) forced on the small table: replica r steps slice r and writes it into every replica."""
    from neuralradiancecaching_b200.dist import FusedExchange
    monkeypatch.setenv("NRC_DP_SHARD", shard)
    T = 16384
    params = Cm.S.mlp_init(1)
    params[20672:] = Cm.S.init_params(7, 323, 0.01)
    tables = Cm.S.table_init(2, 16, T, 2)
    mom = Cm.S.init_params(5, 20995, 0.01)
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=params, tables=tables, lr=(0.05, wd, 0.5), mrhe_lr=(0.1, wd))
    sc.set_mlp_state(np.zeros(20995, np.float32), mom)
    reps, gm, gt = [], [], []
    for r in range(world):
        nrc = N.NeuralRadianceCache(ctx, 0.05, wd, 0.5, seed=0)
        nrc.set_params(params)
        m = N.MRHE(ctx, 0.1, wd, 16, T, 2, 16, 512, Cm.RES_INTENDED, seed=0)
        g = Cm.S.init_params(100 + r, 20995, 0.3)
        t = np.zeros_like(tables)
        t[:2 * T] = Cm.S.init_params(200 + r, 2 * T, 0.2)
        if r == 0:
            g[[3, 4, 20680]] = [np.nan, np.inf, -5000.0]       # scrubbing applies to the SUM
            t[[0, 1]] = [np.nan, -1e6]
        nrc.set_state(delta=g, momentum=mom)
        m.set_tables(tables, t)
        reps.append((nrc, m)); gm.append(g); gt.append(t)
    dps = [FusedExchange(nrc, m, r, world, connect=False) for r, (nrc, m) in enumerate(reps)]
    for d in dps:
        d.connect([x.handle for x in dps], same_process=True)
    if world == 1:
        dps[0]()                                               # the one-call form degenerates to the plain steps
    else:
        for d in dps:
            d.phase(0)
        for d in dps:
            d.phase(1)
    ctx.sync()
    sc.set_mlp_state(_sum_rank_order(gm), mom)
    sc.set_mrhe_delta(_sum_rank_order(gt))
    sc.nrc_step(); sc.mrhe_step()
    op, _, om = sc.get_mlp()
    ot, _ = sc.get_mrhe()
    for nrc, m in reps:
        gp, gd, gmo = nrc.get_params()
        np.testing.assert_array_equal(gp.view(np.uint32), op.view(np.uint32))
        np.testing.assert_array_equal(gmo.view(np.uint32), om.view(np.uint32))
        assert not gd.any()
        tt, td = m.get_tables()
        np.testing.assert_array_equal(tt.view(np.uint32), ot.view(np.uint32))
        assert not td.any()
    # the forward pass after the fused step sees the refreshed weight images (fp32 transposes; bf16 operand images)
    pos, dirs = Cm.rand_queries(512, seed=9)
    want = sc.infer(pos, dirs)
    got = N.infer(reps[-1][0], reps[-1][1], pos, dirs)
    assert float(np.max(np.abs(got - want))) / float(np.max(np.abs(want))) < 1e-3
    for d in dps:
        d.close()
    for nrc, m in reps:
        nrc.Destroy(); m.Destroy()


def test_fused_rejects_unconnected_and_oversized(N, ctx):
    from neuralradiancecaching_b200 import _lib as L
    from neuralradiancecaching_b200.dist import FusedExchange
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0)
    d = FusedExchange(nrc, m, 0, 2, connect=False)
    with pytest.raises(L.NrcError):
        d()                                                    # not connected
    d.close()
    nrc.Destroy(); m.Destroy()


@pytest.mark.parametrize("shard", ["0", "1"])
def test_fused_exchange_with_adam_and_the_fixed_preset(N, ctx, shard, monkeypatch):
    """The opt-in preset through the fused kernel (round 1 refused it): per-level gradient tables (the whole table is exchanged) and
    Adam in both steps.  Three replicas against ONE replica that is handed the rank-ordered sum and stepped by nrc_optimizer_step:
    parameters, tables and all moments bit for bit, over two steps (bias corrections advance)."""
    from neuralradiancecaching_b200.dist import FusedExchange
    monkeypatch.setenv("NRC_DP_SHARD", shard)
    world, T = 3, 4096
    params = Cm.S.mlp_init(1)
    tables = Cm.S.table_init(2, 16, T, 2)

    def replica():
        nrc = N.NeuralRadianceCache(ctx, 0.01, 0.001, 0.9, seed=0)
        nrc.set_params(params)
        m = N.MRHE(ctx, 0.02, 0.001, 16, T, 2, 16, 512, Cm.RES_INTENDED, seed=0)
        m.set_tables(tables)
        m.set_fixed_mode(True)
        nrc.set_adam(True); m.set_adam(True)
        return nrc, m

    reps = [replica() for _ in range(world)]
    ref_nrc, ref_m = replica()
    dps = [FusedExchange(nrc, m, r, world, connect=False) for r, (nrc, m) in enumerate(reps)]
    for d in dps:
        d.connect([x.handle for x in dps], same_process=True)
    for it in range(2):
        gm = [Cm.S.init_params(100 + 10 * it + r, 20995, 0.3) for r in range(world)]
        gt = [Cm.S.init_params(200 + 10 * it + r, 16 * T * 2, 0.2) for r in range(world)]
        for (nrc, m), g, t in zip(reps, gm, gt):
            nrc.set_state(delta=g); m.set_tables(None, t)
        ref_nrc.set_state(delta=_sum_rank_order(gm)); ref_m.set_tables(None, _sum_rank_order(gt))
        N.optimizer_step(ref_nrc, ref_m)
        for d in dps:
            d.phase(0)
        for d in dps:
            d.phase(1)
        ctx.sync()
        rp, _, rm = ref_nrc.get_params()
        rt, _ = ref_m.get_tables()
        rs, _ = ref_nrc.get_adam_state()
        rm1, rm2, _ = ref_m.get_adam_state()
        for nrc, m in reps:
            gp, gd, gmo = nrc.get_params()
            np.testing.assert_array_equal(gp.view(np.uint32), rp.view(np.uint32))
            np.testing.assert_array_equal(gmo.view(np.uint32), rm.view(np.uint32))
            np.testing.assert_array_equal(nrc.get_adam_state()[0].view(np.uint32), rs.view(np.uint32))
            tt, td = m.get_tables()
            np.testing.assert_array_equal(tt.view(np.uint32), rt.view(np.uint32))
            m1, m2, _ = m.get_adam_state()
            np.testing.assert_array_equal(m1.view(np.uint32), rm1.view(np.uint32))
            np.testing.assert_array_equal(m2.view(np.uint32), rm2.view(np.uint32))
            assert not gd.any() and not td.any()
    for d in dps:
        d.close()
    for nrc, m in reps + [(ref_nrc, ref_m)]:
        nrc.Destroy(); m.Destroy()


def test_fused_exchange_fails_closed_when_a_peer_never_arrives(N, ctx, monkeypatch):
    """A rank whose peer never launches gives up after the time-out (NRC_DP_TIMEOUT_MS; default 30 s) and FAILS CLOSED: no step, no
    clear, the mark lands in every rank's flag block, and nrc_sync reports it from then on (the old kernel stepped on
    whatever the late peer's arena held and zeroed gradients the peer had not read)."""
    from neuralradiancecaching_b200 import _lib as L
    from neuralradiancecaching_b200.dist import FusedExchange
    monkeypatch.setenv("NRC_DP_TIMEOUT_MS", "200")
    params = Cm.S.mlp_init(1)
    tables = Cm.S.table_init()
    reps = []
    for r in range(2):
        nrc = N.NeuralRadianceCache(ctx, 0.05, 0.0, 0.5, seed=0)
        nrc.set_params(params)
        m = N.MRHE(ctx, 0.1, 0.0, seed=0)
        g = Cm.S.init_params(100 + r, 20995, 0.3)
        t = np.zeros_like(tables)
        t[:2 * 16384] = Cm.S.init_params(200 + r, 2 * 16384, 0.2)
        nrc.set_state(delta=g)
        m.set_tables(tables, t)
        reps.append((nrc, m, g, t))
    dps = [FusedExchange(nrc, m, r, 2, connect=False) for r, (nrc, m, _, _) in enumerate(reps)]
    for d in dps:
        d.connect([x.handle for x in dps], same_process=True)
    assert not dps[0].timed_out() and not dps[1].timed_out()
    dps[0].phase(2)                       # rank 0's kernel with the real handshakes; "rank 1" never launches
    with pytest.raises(L.NrcError, match="failed closed"):
        ctx.sync()
    assert dps[0].timed_out() and dps[1].timed_out()          # the mark reached the peer's flag block too
    nrc, m, g, t = reps[0]
    gp, gd, _ = nrc.get_params()
    np.testing.assert_array_equal(gp.view(np.uint32), params.view(np.uint32))        # no step
    np.testing.assert_array_equal(gd.view(np.uint32), g.view(np.uint32))             # no clear
    tt, td = m.get_tables()
    np.testing.assert_array_equal(tt.view(np.uint32), tables.view(np.uint32))
    np.testing.assert_array_equal(td.view(np.uint32), t.view(np.uint32))
    with pytest.raises(L.NrcError, match="failed closed"):
        dps[0].phase(2) if False else L.check(L.lib().nrc_sync(ctx.h))
    for d in dps:
        d.close()
    ctx.sync()                            # the context is usable again once the failed exchange objects are gone
    for nrc, m, _, _ in reps:
        nrc.Destroy(); m.Destroy()


def test_fused_exchange_across_gpus():
    """Spawns tests/multi_gpu/check_dp.py on 2 GPUs (skipped on a single-GPU box)."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29611", os.path.join(here, "multi_gpu", "check_dp.py")]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    out = p.stdout.decode()
    assert p.returncode == 0 and "DP-CHECK OK" in out, out[-3000:]
