"""Builds tests/host_emul/walk_host.cpp (the device walk/encode code compiled for the host) and wraps it."""
import ctypes as C
import glob
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libwalk_host.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "walk_host.cpp")
        deps = [src] + glob.glob(os.path.join(_HERE, "../../neuralradiancecaching_b200/csrc/*.cuh")) + \
            [os.path.join(_HERE, "../../include/nrc_math.h")]
        if not os.path.exists(_SO) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps):
            os.makedirs(os.path.dirname(_SO), exist_ok=True)
            gomp = glob.glob("/usr/lib/gcc/x86_64-linux-gnu/*/libgomp.spec")
            flags = ["-B" + os.path.dirname(gomp[0]) + "/"] if gomp else []
            fma = ["-mfma"] if " fma " in open("/proc/cpuinfo").read() else []
            subprocess.check_call([os.environ.get("CXX", "g++")] + flags + ["-O2", "-std=c++17", "-fPIC", "-fopenmp", "-ffp-contract=off"]
                                  + fma + ["-x", "c++", "-shared", "-o", _SO, src])
        _lib = C.CDLL(_SO)
        _lib.emu_density.restype = C.c_float
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f(a):
    return np.ascontiguousarray(a, np.float32)


def set_volume(d):
    d = _f(d)
    lib().emu_set_volume(_p(d), *[C.c_int(s) for s in d.shape])
    return d.shape


def render_walk(cfg, W, H, shape, spp=1):
    cam = cfg["cam"]
    ipv, pos, rnd = _f(cam.invProjView).reshape(16), _f(cam.m_Pos), _f(cfg["random4"][:2])
    cnt = np.zeros((H, W, 4), np.uint32)
    qry = np.zeros((H, W, 7), np.float32)
    lib().emu_render_walk(_p(ipv), _p(pos), _p(rnd), C.c_float(cfg["g"]), C.c_uint(cfg["use_nn"]), C.c_uint(cfg["show_non_nn"]),
                          C.c_uint(W), C.c_uint(H), C.c_uint(spp), C.c_float(cfg["density_factor"]),
                          C.c_int(shape[0]), C.c_int(shape[1]), C.c_int(shape[2]), _p(cnt), _p(qry))
    return cnt, qry


def render_walk_split(cfg, W, H, shape, split_at):
    """the walk as the head + tail launches of trace.cu run it; returns counters, queries and the number of paths handed over"""
    cam = cfg["cam"]
    ipv, pos, rnd = _f(cam.invProjView).reshape(16), _f(cam.m_Pos), _f(cfg["random4"][:2])
    cnt = np.zeros((H, W, 4), np.uint32)
    qry = np.zeros((H, W, 7), np.float32)
    n = C.c_uint(0)
    lib().emu_render_walk_split(_p(ipv), _p(pos), _p(rnd), C.c_float(cfg["g"]), C.c_uint(cfg["use_nn"]), C.c_uint(cfg["show_non_nn"]),
                                C.c_uint(W), C.c_uint(H), C.c_int(split_at), C.c_float(cfg["density_factor"]),
                                C.c_int(shape[0]), C.c_int(shape[1]), C.c_int(shape[2]), _p(cnt), _p(qry), C.byref(n))
    return cnt, qry, n.value


def render_walk_cost(cfg, W, H, shape):
    """divergence study (scripts/walk_divergence.py): per pixel {walk steps, SDF evaluations of the walk, of the camera ray, lit
    vertices} and the SDF evaluations of every walk step"""
    cam = cfg["cam"]
    ipv, pos, rnd = _f(cam.invProjView).reshape(16), _f(cam.m_Pos), _f(cfg["random4"][:2])
    cost = np.zeros((H, W, 4), np.uint32)
    per = np.zeros((H, W, 32), np.uint8)
    lib().emu_render_walk_cost(_p(ipv), _p(pos), _p(rnd), C.c_float(cfg["g"]), C.c_uint(cfg["use_nn"]), C.c_uint(W), C.c_uint(H),
                               C.c_float(cfg["density_factor"]), C.c_int(shape[0]), C.c_int(shape[1]), C.c_int(shape[2]), _p(cost), _p(per))
    return cost, per


def train_walk(cfg, tw, th, shape, env_light=False):
    cam = cfg["cam"]
    ipv, pos, rnd = _f(cam.invProjView).reshape(16), _f(cam.m_Pos), _f(cfg["random4"][:2])
    rec = np.zeros((tw * th, 6), np.float32)
    cnt = np.zeros((tw * th, 3), np.uint32)
    lib().emu_train_walk(_p(ipv), _p(pos), _p(rnd), C.c_float(cfg["g"]), C.c_uint(tw), C.c_uint(th), C.c_float(cfg["density_factor"]),
                         C.c_int(shape[0]), C.c_int(shape[1]), C.c_int(shape[2]), _p(rec), _p(cnt), C.c_int(1 if env_light else 0))
    return rec, cnt


def encode(pos, tables, L, T, res):
    pos, tables = _f(pos), _f(tables)
    res = np.ascontiguousarray(res, np.uint32)
    n = pos.shape[0]
    feat = np.zeros((n, 32), np.float32)
    idx = np.zeros((n, 128), np.uint32)
    lib().emu_encode(_p(pos), C.c_long(n), _p(tables), C.c_uint(L), C.c_uint(T), _p(res), _p(feat), _p(idx))
    return feat, idx


def oneblob(d):
    d = _f(d)
    feat = np.zeros((d.shape[0], 32), np.float32)
    lib().emu_oneblob(_p(d), C.c_long(d.shape[0]), _p(feat))
    return feat


def density_at(p, density_factor, shape):
    p = _f(p)
    return float(lib().emu_density(_p(p), C.c_float(density_factor), C.c_int(shape[0]), C.c_int(shape[1]), C.c_int(shape[2])))
