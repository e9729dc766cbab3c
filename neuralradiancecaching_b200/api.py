"""Host-side mirror of the reference's classes over the C-ABI (names and argument meaning follow
include/engine/graphics/{NeuralRadianceCache,MRHE,HdrEnvMap,DirLight,PointLight,Camera}.hpp,
include/engine/objects/VolumeData.hpp and include/engine/graphics/renderer/NrcHpmRenderer.hpp of the reference;
Vulkan handle parameters are dropped).  Arrays may be numpy arrays (host memory) or torch CUDA tensors
(device memory): the library detects which.
"""
import ctypes as C

import numpy as np

from . import _lib as L


def _f32(a):
    """numpy float32 contiguous view (host) or the object itself when it is a CUDA tensor."""
    if hasattr(a, "data_ptr"):
        assert a.is_contiguous() and str(a.dtype) == "torch.float32", "device arrays must be contiguous float32"
        return a
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return a.ctypes.data_as(C.c_void_p)


def _out(shape, dtype, like):
    """Output buffer on the same side as `like` (torch CUDA tensor -> torch, else numpy)."""
    if hasattr(like, "data_ptr") and like.is_cuda:
        import torch
        tdt = {np.float32: torch.float32, np.uint32: torch.int32}[dtype]
        return torch.empty(shape, dtype=tdt, device=like.device)
    return np.empty(shape, dtype)


class Context:
    """Replaces VulkanAPI::Init/Shutdown (src/VulkanAPI.cpp:35-65): one device, one stream."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        L.check(L.lib().nrc_ctx_create(C.c_int(device), C.byref(self.h)))
        self.device = device

    def sync(self):
        L.check(L.lib().nrc_sync(self.h))

    @property
    def stream(self):
        return L.lib().nrc_ctx_stream(self.h)

    @property
    def sm_count(self):
        return int(L.lib().nrc_ctx_sm_count(self.h))

    def launch_count(self):
        return int(L.lib().nrc_ctx_launch_count(self.h))

    def profile(self, on):
        """Bracket every kernel launch with CUDA events on its stream (per-kernel device time)."""
        L.check(L.lib().nrc_ctx_profile_enable(self.h, C.c_int(1 if on else 0)))

    def profile_read(self, reset=True):
        """{kernel name: (total ms, launches)} accumulated since the last reset."""
        k = int(L.lib().nrc_kernel_count())
        ms = (C.c_double * k)()
        cnt = (C.c_uint64 * k)()
        L.check(L.lib().nrc_ctx_profile_read(self.h, ms, cnt, C.c_int(1 if reset else 0)))
        return {L.lib().nrc_kernel_name(i).decode(): (ms[i], int(cnt[i])) for i in range(k) if cnt[i]}

    def profile_timeline(self, max_records=4096):
        """[(kernel name, start ms, end ms)] of the launches since the last read, relative to the first."""
        kid = (C.c_int * max_records)()
        s = (C.c_float * max_records)()
        e = (C.c_float * max_records)()
        n = C.c_int()
        L.check(L.lib().nrc_ctx_profile_timeline(self.h, C.c_int(max_records), kid, s, e, C.byref(n)))
        return [(L.lib().nrc_kernel_name(kid[i]).decode(), s[i], e[i]) for i in range(n.value)]

    def set_precision(self, p):
        L.check(L.lib().nrc_ctx_set_precision(self.h, C.c_int(p)))

    def set_infer_mode(self, split):
        """NRC_INFER_FUSED (k_infer_tc) or NRC_INFER_SPLIT (k_encode_tiles -> k_mlp_stream) for the bf16 tensor-core inference."""
        L.check(L.lib().nrc_ctx_set_infer_mode(self.h, C.c_int(1 if split else 0)))

    def destroy(self):
        if self.h:
            L.lib().nrc_ctx_destroy(self.h)
            self.h = C.c_void_p()


class NeuralRadianceCache:
    """en::NeuralRadianceCache(learningRate, weightDecay, beta1) -- NeuralRadianceCache.hpp:16-30."""

    class StatsData:
        def __init__(self, mse_loss):
            self.mseLoss = mse_loss

    def __init__(self, ctx, learning_rate, weight_decay, beta1, seed=1):
        self.ctx = ctx
        self.h = C.c_void_p()
        L.check(L.lib().nrc_mlp_create(ctx.h, C.c_float(learning_rate), C.c_float(weight_decay), C.c_float(beta1),
                                       C.c_uint64(seed), C.byref(self.h)))
        self.adam = False

    def Destroy(self):
        if self.h:
            L.lib().nrc_mlp_destroy(self.h)
            self.h = C.c_void_p()

    def set_adam(self, on, beta2=0.999, eps=1e-8):
        """opt-in fixed preset: second moment + bias correction in nrc-step (resets moments and step count)"""
        L.check(L.lib().nrc_mlp_set_adam(self.h, C.c_int(1 if on else 0), C.c_float(beta2), C.c_float(eps)))
        self.adam = bool(on)

    def get_adam_state(self):
        """(second moment [20995], step count); only while Adam is on"""
        v = np.empty(L.PARAM_COUNT, np.float32)
        n = C.c_uint32()
        L.check(L.lib().nrc_mlp_get_adam_state(self.h, _ptr(v), C.byref(n)))
        return v, int(n.value)

    def set_adam_state(self, second, steps):
        L.check(L.lib().nrc_mlp_set_adam_state(self.h, _ptr(_f32(second)), C.c_uint32(steps)))

    def ResetStats(self):
        L.check(L.lib().nrc_mlp_reset_stats(self.h))

    def GetStats(self):
        v = C.c_float()
        L.check(L.lib().nrc_mlp_get_stats(self.h, C.byref(v)))
        return NeuralRadianceCache.StatsData(v.value)

    def get_stats_async(self, out):
        """Queue the copy of mseLoss into out[0] (a page-locked float32 array); valid after readback_wait()/sync()."""
        L.check(L.lib().nrc_mlp_get_stats_async(self.h, _ptr(out)))

    def PrintWeights(self):
        p = self.get_params()[0]
        for i in range(6):
            n = 4096 if i < 5 else 192
            w = p[i * 4096:i * 4096 + n]
            print("Weights %d: [%s]" % (i, ", ".join("%f" % x for x in w)))

    # checkpoint / parity surface
    def set_config(self, lr, wd, beta1):
        L.check(L.lib().nrc_mlp_set_config(self.h, C.c_float(lr), C.c_float(wd), C.c_float(beta1)))

    def set_params(self, params):
        p = _f32(params)
        L.check(L.lib().nrc_mlp_set_params(self.h, _ptr(p)))

    def set_state(self, delta=None, momentum=None):
        d = None if delta is None else _f32(delta)
        m = None if momentum is None else _f32(momentum)
        L.check(L.lib().nrc_mlp_set_state(self.h, _ptr(d), _ptr(m)))

    def get_params(self):
        p = np.empty(L.PARAM_COUNT, np.float32)
        d = np.empty(L.PARAM_COUNT, np.float32)
        m = np.empty(L.PARAM_COUNT, np.float32)
        L.check(L.lib().nrc_mlp_get_params(self.h, _ptr(p), _ptr(d), _ptr(m)))
        return p, d, m

    def grad_ptr(self):
        return L.lib().nrc_mlp_grad_ptr(self.h), int(L.lib().nrc_mlp_grad_count(self.h))


class MRHE:
    """en::MRHE(learningRate, weightDecay) -- MRHE.hpp:14-22; the constants of src/MRHE.cpp:83-87 become arguments."""

    def __init__(self, ctx, learning_rate, weight_decay, level_count=16, hash_table_size=16384, feature_count=2,
                 min_res=16, max_res=512, resolutions=None, seed=2):
        self.ctx = ctx
        self.L, self.T, self.F = level_count, hash_table_size, feature_count
        self.h = C.c_void_p()
        res = None if resolutions is None else np.ascontiguousarray(resolutions, dtype=np.uint32)
        assert res is None or res.size == 16
        L.check(L.lib().nrc_mrhe_create(ctx.h, C.c_float(learning_rate), C.c_float(weight_decay), C.c_uint32(level_count),
                                        C.c_uint32(hash_table_size), C.c_uint32(feature_count), C.c_uint32(min_res),
                                        C.c_uint32(max_res), _ptr(res), C.c_uint64(seed), C.byref(self.h)))
        self.adam = False

    def Destroy(self):
        if self.h:
            L.lib().nrc_mrhe_destroy(self.h)
            self.h = C.c_void_p()

    def GetHashTableSize(self):
        return int(L.lib().nrc_mrhe_table_bytes(self.h))

    def PrintHashTables(self):
        t = self.get_tables()[0]
        print("HashTables: [%s]" % ", ".join("%f" % x for x in t[:256]))

    @property
    def count(self):
        return self.L * self.T * self.F

    def set_config(self, lr, wd):
        L.check(L.lib().nrc_mrhe_set_config(self.h, C.c_float(lr), C.c_float(wd)))

    def set_fixed_mode(self, on):
        """opt-in fixed preset: grid-space lerp factors on their own axes, per-feature errors, per-level gradient tables"""
        L.check(L.lib().nrc_mrhe_set_fixed_mode(self.h, C.c_int(1 if on else 0)))

    def set_adam(self, on, beta1=0.9, beta2=0.999, eps=1e-8):
        """opt-in fixed preset: Adam in mrhe-step (resets moments and step count)"""
        L.check(L.lib().nrc_mrhe_set_adam(self.h, C.c_int(1 if on else 0), C.c_float(beta1), C.c_float(beta2), C.c_float(eps)))
        self.adam = bool(on)

    def get_adam_state(self):
        """(first moment, second moment [L*T*F each], step count); only while Adam is on"""
        m1, m2 = np.empty(self.count, np.float32), np.empty(self.count, np.float32)
        n = C.c_uint32()
        L.check(L.lib().nrc_mrhe_get_adam_state(self.h, _ptr(m1), _ptr(m2), C.byref(n)))
        return m1, m2, int(n.value)

    def set_adam_state(self, m1, m2, steps):
        L.check(L.lib().nrc_mrhe_set_adam_state(self.h, _ptr(_f32(m1)), _ptr(_f32(m2)), C.c_uint32(steps)))

    def set_tables(self, tables=None, delta=None):
        t = None if tables is None else _f32(tables)
        d = None if delta is None else _f32(delta)
        L.check(L.lib().nrc_mrhe_set_tables(self.h, _ptr(t), _ptr(d)))

    def get_tables(self):
        t = np.empty(self.count, np.float32)
        d = np.empty(self.count, np.float32)
        L.check(L.lib().nrc_mrhe_get_tables(self.h, _ptr(t), _ptr(d)))
        return t, d

    def grad_ptr(self):
        return L.lib().nrc_mrhe_grad_ptr(self.h), int(L.lib().nrc_mrhe_grad_extent(self.h))


def mrhe_resolutions(min_res=16, max_res=512, level_count=16, mode=L.RES_INTENDED):
    out = (C.c_uint32 * 16)()
    L.check(L.lib().nrc_mrhe_resolutions(C.c_uint32(min_res), C.c_uint32(max_res), C.c_uint32(level_count), C.c_int(mode), out))
    return np.array(list(out), dtype=np.uint32)


# ---- stand-alone operators ----
def encode(mrhe, pos, want_idx=True):
    pos = _f32(pos)
    n = pos.shape[0]
    feat = _out((n, 32), np.float32, pos)
    idx = _out((n, 128), np.uint32, pos) if want_idx else None
    L.check(L.lib().nrc_encode(mrhe.h, _ptr(pos), C.c_size_t(n), _ptr(feat), _ptr(idx)))
    return feat, idx


def encode_dir(ctx, d):
    d = _f32(d)
    n = d.shape[0]
    feat = _out((n, 32), np.float32, d)
    L.check(L.lib().nrc_encode_dir(ctx.h, _ptr(d), C.c_size_t(n), _ptr(feat)))
    return feat


def infer(nrc, mrhe, pos, d, out=None):
    pos, d = _f32(pos), _f32(d)
    n = pos.shape[0]
    rgb = _out((n, 3), np.float32, pos) if out is None else out
    L.check(L.lib().nrc_infer(nrc.h, mrhe.h, _ptr(pos), _ptr(d), C.c_size_t(n), _ptr(rgb)))
    return rgb


def feature_tile_bytes(n):
    return int(L.lib().nrc_feature_tile_bytes(C.c_size_t(n)))


def encode_tiles(mrhe, pos, d, out=None):
    """EncodeRay into bf16 operand-image tiles (device memory: a torch CUDA uint8 tensor of feature_tile_bytes(n))."""
    import torch
    pos, d = _f32(pos), _f32(d)
    n = pos.shape[0]
    tiles = torch.empty(feature_tile_bytes(n), dtype=torch.uint8, device="cuda") if out is None else out
    L.check(L.lib().nrc_encode_tiles(mrhe.h, _ptr(pos), _ptr(d), C.c_size_t(n), _ptr(tiles)))
    return tiles


def mlp_forward_tiles(nrc, tiles, n, out=None, like=None):
    """The six layers from bf16 operand-image tiles (k_mlp_stream): rgb[n][3], numpy unless `out`/`like` is a CUDA tensor."""
    rgb = out if out is not None else _out((n, 3), np.float32, like)
    L.check(L.lib().nrc_mlp_forward_tiles(nrc.h, _ptr(tiles), C.c_size_t(n), _ptr(rgb)))
    return rgb


def train_accumulate(nrc, mrhe, pos, d, target, loss_scale, want_pred=False):
    pos, d, target = _f32(pos), _f32(d), _f32(target)
    n = pos.shape[0]
    pred = _out((n, 3), np.float32, pos) if want_pred else None
    L.check(L.lib().nrc_train_accumulate(nrc.h, mrhe.h, _ptr(pos), _ptr(d), _ptr(target), C.c_size_t(n), C.c_float(loss_scale),
                                         _ptr(pred)))
    return pred


def optimizer_step(nrc, mrhe):
    L.check(L.lib().nrc_optimizer_step(None if nrc is None else nrc.h, None if mrhe is None else mrhe.h))
