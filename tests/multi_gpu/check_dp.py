"""Run under torchrun on W >= 2 GPUs: the fused peer-memory exchange (csrc/dp.cu) against NCCL all-reduce + step kernels
from the same per-rank gradients.  W == 2: bit-identical (a+b has one order); W > 2: replicas of the fused path are
bit-identical to EACH OTHER and to a host sum in rank order; the NCCL result agrees to fp32 rounding.
Prints 'DP-CHECK OK' on rank 0; any mismatch raises."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import common as Cm  # noqa: E402
import neuralradiancecaching_b200 as N  # noqa: E402
from neuralradiancecaching_b200.dist import FusedExchange, GradientExchange  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = N.Context(local)
    T = 16384
    params = Cm.S.mlp_init(1)
    tables = Cm.S.table_init(2, 16, T, 2)
    mom = Cm.S.init_params(5, 20995, 0.01)

    def replica():
        nrc = N.NeuralRadianceCache(ctx, 0.05, 0.01, 0.5, seed=0)
        nrc.set_params(params)
        m = N.MRHE(ctx, 0.1, 0.01, 16, T, 2, 16, 512, Cm.RES_INTENDED, seed=0)
        m.set_tables(tables)
        return nrc, m

    a_nrc, a_m = replica()     # NCCL path
    b_nrc, b_m = replica()     # fused path
    nccl = GradientExchange(ctx, a_nrc, a_m, dev)
    fused = FusedExchange(b_nrc, b_m, rank, world, dev)
    for it in range(5):        # several frames: the flag epochs advance, the arenas are cleared and refilled
        g = Cm.S.init_params(1000 * it + rank, 20995, 0.3)
        t = np.zeros_like(tables)
        t[:2 * T] = Cm.S.init_params(7000 * it + rank, 2 * T, 0.2)
        if it == 0:
            mm = mom
        else:
            mm = None
        for nrc, m in ((a_nrc, a_m), (b_nrc, b_m)):
            nrc.set_state(delta=g, momentum=mm)
            m.set_tables(None, t)
        ctx.sync()
        nccl()
        N.optimizer_step(a_nrc, a_m)
        fused()
        ctx.sync()
        ap, ad, am = a_nrc.get_params()
        bp, bd, bm = b_nrc.get_params()
        at, atd = a_m.get_tables()
        bt, btd = b_m.get_tables()
        assert not bd.any() and not btd.any(), "fused path left gradients behind"
        if world == 2:
            np.testing.assert_array_equal(ap.view(np.uint32), bp.view(np.uint32))
            np.testing.assert_array_equal(am.view(np.uint32), bm.view(np.uint32))
            np.testing.assert_array_equal(at.view(np.uint32), bt.view(np.uint32))
        else:
            np.testing.assert_allclose(ap, bp, rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(at, bt, rtol=1e-5, atol=1e-6)
        # every replica of the fused path holds the same bits
        mine = torch.from_numpy(np.concatenate([bp, bt]).view(np.int32).copy()).to(dev)
        lo, hi = mine.clone(), mine.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert bool((lo == hi).all()), "replicas diverged"
    assert not fused.timed_out()
    dist.barrier()
    fused.close()
    if rank == 0:
        print("DP-CHECK OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
