"""-m gpu: nrc_train_accumulate_static -- the allocation-free variant of nrc_train_accumulate (static __device__ scratch, no
cudaMalloc; no counterpart in the reference) -- against the existing entry point,
through the C-ABI, on two calls in a row.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import common as Cm

pytestmark = pytest.mark.gpu

REL = 1e-3   # the fp32 path's bar (tests/test_gpu_ops.py): max abs error over the tensor's largest magnitude


def _rel(a, b):
    return float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64))) / max(float(np.max(np.abs(b))), 1e-30))


def test_static_scratch_train_accumulate_matches_the_allocating_one():
    import torch
    import neuralradiancecaching_b200 as N
    from neuralradiancecaching_b200 import _lib as L
    lib = L.lib()
    ctx = N.Context(0)
    params, tables = Cm.S.mlp_init(1), Cm.S.table_init()
    params[20672:] = Cm.S.init_params(7, 323, 0.01)
    cap = int(lib.nrc_train_accumulate_static_max_samples())
    assert cap == 8192
    dev = torch.device("cuda", 0)

    def fresh():
        nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
        nrc.set_params(params)
        m = N.MRHE(ctx, 0.01, 0.0, seed=0)
        m.set_tables(tables)
        return nrc, m

    ref_nrc, ref_m = fresh()
    got_nrc, got_m = fresh()
    # two calls in a row on the same handles: different sizes, the second smaller than the first (stale scratch rows past n must not
    # leak), gradients accumulate over both like the existing entry point's
    for call, n in enumerate((5000, 777)):
        pos, d = Cm.rand_queries(n, seed=900 + call)
        target = (np.random.Generator(np.random.PCG64(5 + call)).random((n, 3)) * 3.0).astype(np.float32)
        target[3] = [2000.0, 0.5, 0.1]      # the 1024 clamp
        pos[9] = [np.nan, 0.0, 0.0]         # the NaN guard
        scale = 1.0 / n
        want_pred = N.train_accumulate(ref_nrc, ref_m, pos, d, target, scale, want_pred=True)
        dpos, dd, dt = (torch.from_numpy(a).to(dev).contiguous() for a in (pos, d, target))
        dpred = torch.full((n, 3), -1.0, dtype=torch.float32, device=dev)
        ctx.sync()
        torch.cuda.synchronize()
        L.check(lib.nrc_train_accumulate_static(got_nrc.h, got_m.h, C.c_void_p(dpos.data_ptr()), C.c_void_p(dd.data_ptr()),
                                                C.c_void_p(dt.data_ptr()), C.c_size_t(n), C.c_float(scale), C.c_void_p(dpred.data_ptr())))
        ctx.sync()
        got_pred = dpred.cpu().numpy()
        assert np.isfinite(got_pred).all()
        assert _rel(got_pred, want_pred) < REL, call
        gd, wd = got_nrc.get_params()[1], ref_nrc.get_params()[1]
        assert np.isfinite(gd).all() and _rel(gd, wd) < REL, (call, _rel(gd, wd))
        gt, wt = got_m.get_tables()[1], ref_m.get_tables()[1]
        assert np.count_nonzero(gt[2 * 16384:]) == 0
        assert _rel(gt[:2 * 16384], wt[:2 * 16384]) < REL, (call, _rel(gt[:2 * 16384], wt[:2 * 16384]))
        assert abs(got_nrc.GetStats().mseLoss - ref_nrc.GetStats().mseLoss) <= REL * abs(ref_nrc.GetStats().mseLoss)
    # limits: more samples than the static scratch holds, and host pointers (staging them would allocate), are refused with a clear error
    pos, d = Cm.rand_queries(16, seed=1)
    st = lib.nrc_train_accumulate_static(got_nrc.h, got_m.h, C.c_void_p(dpos.data_ptr()), C.c_void_p(dd.data_ptr()), C.c_void_p(dt.data_ptr()),
                                         C.c_size_t(cap + 1), C.c_float(1.0), None)
    assert st == 1 and b"8192" in lib.nrc_last_error()
    st = lib.nrc_train_accumulate_static(got_nrc.h, got_m.h, pos.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p),
                                         pos.ctypes.data_as(C.c_void_p), C.c_size_t(16), C.c_float(1.0), None)
    assert st == 1 and b"device memory" in lib.nrc_last_error()
    for o in (ref_nrc, ref_m, got_nrc, got_m):
        o.Destroy()
    ctx.destroy()
