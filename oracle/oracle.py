"""ctypes front end of the CPU oracle (oracle/_build/liboracle.so).

TEST INFRASTRUCTURE -- parity unpinned (see oracle/nrc_oracle.hpp).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

PARAM_COUNT = 20995
WEIGHT_OFFSET = (0, 4096, 8192, 12288, 16384, 20480)
BIAS_OFFSET = (20672, 20736, 20800, 20864, 20928, 20992)


def build(force=False):
    """Compile the oracle with its Makefile (g++ only; seconds)."""
    src_there = os.path.exists(os.path.join(_HERE, "Makefile"))
    if force or not os.path.exists(_SO) or src_there:   # make is a no-op when the library is up to date
        subprocess.check_call(["make", "-C", _HERE] + (["-B"] if force else []),
                              stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_scene_create.restype = C.c_void_p
        _lib.orc_get_mse_loss.restype = C.c_float
        _lib.orc_hash.restype = C.c_uint
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """One reference scene: the seven descriptor sets of the hot path as host arrays."""

    def __init__(self):
        self.l = lib()
        self.h = C.c_void_p(self.l.orc_scene_create())
        self.table_floats = 0

    def __del__(self):
        try:
            self.l.orc_scene_destroy(self.h)
        except Exception:
            pass

    # ---- state ----
    def set_camera(self, proj_view, inv_proj_view, pos):
        pv, ipv, p = _f(proj_view).reshape(16), _f(inv_proj_view).reshape(16), _f(pos)
        self.l.orc_set_camera(self.h, _p(pv), _p(ipv), _p(p))

    def set_volume(self, density_xyz):
        d = _f(density_xyz)
        nx, ny, nz = d.shape
        self.l.orc_set_volume(self.h, _p(d), C.c_int(nx), C.c_int(ny), C.c_int(nz))

    def set_volume_params(self, random4=(0, 0, 0, 0), use_nn=0, show_non_nn=0, density_factor=0.4, g=0.7,
                          no_nn_spp=1, with_nn_spp=1):
        r = _f(random4)
        self.l.orc_set_volume_params(self.h, _p(r), C.c_uint(use_nn), C.c_uint(show_non_nn), C.c_float(density_factor),
                                     C.c_float(g), C.c_int(no_nn_spp), C.c_int(with_nn_spp))

    def set_dir_light(self, color, zenith, direction, azimuth, strength):
        c, d = _f(color), _f(direction)
        self.l.orc_set_dir_light(self.h, _p(c), C.c_float(zenith), _p(d), C.c_float(azimuth), C.c_float(strength))

    def set_point_light(self, pos, strength, color):
        p, c = _f(pos), _f(color)
        self.l.orc_set_point_light(self.h, _p(p), C.c_float(strength), _p(c))

    def set_envmap(self, rgba, inv_cdf_x=None, inv_cdf_y=None, direct_strength=1.0, hpm_strength=8.0):
        rgba = _f(rgba)
        h, w = rgba.shape[:2]
        cx = None if inv_cdf_x is None else _f(inv_cdf_x)
        cy = None if inv_cdf_y is None else _f(inv_cdf_y)
        self.l.orc_set_envmap(self.h, C.c_int(w), C.c_int(h), _p(rgba), _p(cx), _p(cy), C.c_float(direct_strength),
                              C.c_float(hpm_strength))

    def set_train_env_light(self, on):
        """opt-in: training targets include the env-map gather (the term commented out at train.comp:1118)"""
        self.l.orc_set_train_env_light(self.h, C.c_int(1 if on else 0))

    def set_env_cdf_half(self, on):
        """opt-in: 4 phase-sampled + 4 CDF-importance-sampled env directions per vertex (fwd.frag:715 as commented out there)"""
        self.l.orc_set_env_cdf_half(self.h, C.c_int(1 if on else 0))

    def set_mrhe_fixed(self, on):
        """opt-in "fixed" preset: grid-space lerp factors on their own axes, per-feature errors, per-level gradient tables"""
        self.l.orc_set_mrhe_fixed(self.h, C.c_int(1 if on else 0))

    def set_adam(self, on, beta2=0.999, eps=1e-8, mrhe_beta1=0.9):
        """opt-in "fixed" preset: Adam (second moment + bias correction) in both step shaders; resets the moments"""
        self.l.orc_set_adam(self.h, C.c_int(1 if on else 0), C.c_float(beta2), C.c_float(eps), C.c_float(mrhe_beta1))

    def set_mlp(self, params, lr=0.001, wd=0.0, beta1=0.5):
        p = None if params is None else _f(params)
        assert p is None or p.size == PARAM_COUNT
        self.l.orc_set_mlp(self.h, _p(p), C.c_float(lr), C.c_float(wd), C.c_float(beta1))

    def set_mlp_state(self, delta=None, momentum=None):
        d = None if delta is None else _f(delta)
        m = None if momentum is None else _f(momentum)
        self.l.orc_set_mlp_state(self.h, _p(d), _p(m))

    def get_mlp(self):
        p = np.empty(PARAM_COUNT, np.float32)
        d = np.empty(PARAM_COUNT, np.float32)
        m = np.empty(PARAM_COUNT, np.float32)
        self.l.orc_get_mlp(self.h, _p(p), _p(d), _p(m))
        return p, d, m

    def set_mrhe(self, tables, resolutions, lr=0.01, wd=0.0, L=16, T=16384, F=2, min_res=16, max_res=512):
        res = np.ascontiguousarray(resolutions, dtype=np.uint32)
        assert res.size == 16
        t = None if tables is None else _f(tables)
        self.table_floats = L * T * F
        assert t is None or t.size == self.table_floats
        self.l.orc_set_mrhe(self.h, C.c_float(lr), C.c_float(wd), C.c_uint(L), C.c_uint(T), C.c_uint(F), C.c_uint(min_res),
                            C.c_uint(max_res), _p(res), _p(t))

    def set_mrhe_delta(self, delta):
        d = _f(delta)
        assert d.size == self.table_floats
        self.l.orc_set_mrhe_delta(self.h, _p(d))

    def get_mrhe(self):
        t = np.empty(self.table_floats, np.float32)
        d = np.empty(self.table_floats, np.float32)
        self.l.orc_get_mrhe(self.h, _p(t), _p(d))
        return t, d

    def set_train_size(self, w, h):
        self.l.orc_set_train_size(self.h, C.c_uint(w), C.c_uint(h))

    def set_loss_scale(self, wf, hf):
        self.l.orc_set_loss_scale(self.h, C.c_float(wf), C.c_float(hf))

    def mse_loss(self):
        return float(self.l.orc_get_mse_loss(self.h))

    def reset_stats(self):
        self.l.orc_reset_stats(self.h)

    # ---- ops ----
    def encode_pos(self, pos):
        pos = _f(pos).reshape(-1, 3)
        n = pos.shape[0]
        feat = np.empty((n, 32), np.float32)
        idx = np.empty((n, 128), np.uint32)
        lerp = np.empty((n, 16, 3), np.float32)
        self.l.orc_encode_pos(self.h, _p(pos), C.c_long(n), _p(feat), _p(idx), _p(lerp))
        return feat, idx, lerp

    def encode_dir(self, d):
        d = _f(d).reshape(-1, 3)
        n = d.shape[0]
        feat = np.empty((n, 32), np.float32)
        self.l.orc_encode_dir(self.h, _p(d), C.c_long(n), _p(feat))
        return feat

    def density(self, pos):
        pos = _f(pos).reshape(-1, 3)
        out = np.empty(pos.shape[0], np.float32)
        self.l.orc_density(self.h, _p(pos), C.c_long(pos.shape[0]), _p(out))
        return out

    def infer(self, pos, d):
        pos, d = _f(pos).reshape(-1, 3), _f(d).reshape(-1, 3)
        n = pos.shape[0]
        rgb = np.empty((n, 3), np.float32)
        self.l.orc_infer(self.h, _p(pos), _p(d), C.c_long(n), _p(rgb))
        return rgb

    def train_accumulate(self, pos, d, target):
        pos, d, target = _f(pos).reshape(-1, 3), _f(d).reshape(-1, 3), _f(target).reshape(-1, 3)
        n = pos.shape[0]
        pred = np.empty((n, 3), np.float32)
        dropped = np.empty(n, np.uint8)
        self.l.orc_train_accumulate(self.h, _p(pos), _p(d), _p(target), C.c_long(n), _p(pred), _p(dropped))
        return pred, dropped

    def nrc_step(self):
        self.l.orc_nrc_step(self.h)

    def mrhe_step(self):
        self.l.orc_mrhe_step(self.h)

    def render(self, w, h, y0=0, y1=None, want_counters=True, want_query=True):
        y1 = h if y1 is None else y1
        rgba = np.zeros((h, w, 4), np.float32)
        cnt = np.zeros((h, w, 5), np.uint32) if want_counters else None
        qry = np.zeros((h, w, 8), np.float32) if want_query else None
        self.l.orc_render(self.h, C.c_uint(w), C.c_uint(h), C.c_uint(y0), C.c_uint(y1), _p(rgba), _p(cnt), _p(qry))
        return rgba, cnt, qry

    def train_trace(self, tw, th, i0=0, i1=None):
        i1 = tw * th if i1 is None else i1
        n = i1 - i0
        rec = np.zeros((n, 9), np.float32)
        cnt = np.zeros((n, 5), np.uint32)
        self.l.orc_train_trace(self.h, C.c_uint(tw), C.c_uint(th), C.c_long(i0), C.c_long(i1), _p(rec), _p(cnt))
        return rec, cnt

    def frame(self, w, h, tw, th):
        rgba = np.zeros((h, w, 4), np.float32)
        self.l.orc_frame(self.h, C.c_uint(w), C.c_uint(h), C.c_uint(tw), C.c_uint(th), _p(rgba))
        return rgba


def mrhe_resolutions(min_res=16, max_res=512, levels=16):
    out = np.zeros(16, np.uint32)
    lib().orc_mrhe_resolutions(C.c_uint(min_res), C.c_uint(max_res), C.c_uint(levels), _p(out))
    return out


def hdr4f_to_cdf(rgba):
    rgba = _f(rgba)
    h, w = rgba.shape[:2]
    cx = np.empty((h, w), np.float32)
    cy = np.empty(h, np.float32)
    lib().orc_hdr4f_to_cdf(_p(rgba), C.c_long(w), C.c_long(h), _p(cx), _p(cy))
    return cx, cy


_FN = {"sin": 0, "cos": 1, "exp": 2, "acos": 3, "asin": 4, "atan2": 5, "pow15": 6, "rand": 7}


def math_eval(fn, x, y=None):
    x = _f(x).ravel()
    y = np.zeros_like(x) if y is None else _f(y).ravel()
    out = np.empty_like(x)
    lib().orc_math_eval(C.c_int(_FN[fn]), _p(x), _p(y), C.c_long(x.size), _p(out))
    return out


def hash3(x, y, z, T):
    return int(lib().orc_hash(C.c_uint(x), C.c_uint(y), C.c_uint(z), C.c_uint(T)))


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(int(n)))


def num_threads():
    return int(lib().orc_num_threads())
