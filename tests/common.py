"""Shared fixtures: one deterministic scene, set up identically for the CPU oracle and the CUDA library."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from neuralradiancecaching_b200 import scene as S          # noqa: E402
from neuralradiancecaching_b200.objects import Camera, vec_from_angles   # noqa: E402

SKY = np.array([62.5, 42.5, 76.5], np.float32)
RES_INTENDED = np.array([16, 20, 25, 32, 40, 50, 64, 80, 101, 128, 161, 203, 256, 322, 406, 512], np.uint32)
RES_SHIPPED = np.array([20, 50, 128, 322] + [0] * 12, np.uint32)

_cache = {}


def density():
    if "d" not in _cache:
        _cache["d"] = S.cloud_density()
    return _cache["d"]


def envmap():
    if "e" not in _cache:
        _cache["e"] = S.sky_envmap(256, 128)
    return _cache["e"]


def camera(aspect=1.0):
    # src/main.cpp:142-149
    return Camera((0.0, 0.0, -64.0), (0.0, 0.0, 1.0), (0.0, 1.0, 0.0), aspect, np.radians(60.0), 0.1, 100.0)


DIR_ZENITH, DIR_AZIMUTH = -1.57, 0.0     # src/main.cpp:151


def scene_cfg(random4=(0.37, 0.61, 0.13, 0.89), use_nn=1, dir_strength=1.0, pt_strength=1.0, aspect=1.0):
    cam = camera(aspect)
    return dict(cam=cam, random4=np.asarray(random4, np.float32), use_nn=use_nn, show_non_nn=0, density_factor=0.4, g=0.7,
                dir_dir=vec_from_angles(DIR_ZENITH, DIR_AZIMUTH), dir_strength=dir_strength,
                pt_pos=np.zeros(3, np.float32), pt_color=np.ones(3, np.float32), pt_strength=pt_strength,
                env_direct=1.0, env_hpm=8.0)


def oracle_scene(cfg, params=None, tables=None, res=RES_INTENDED, T=16384, lr=(0.001, 0.0, 0.5), mrhe_lr=(0.01, 0.0)):
    from oracle import oracle as O
    sc = O.OracleScene()
    cam = cfg["cam"]
    sc.set_camera(cam.projView, cam.invProjView, cam.m_Pos)
    sc.set_volume(density())
    sc.set_volume_params(cfg["random4"], cfg["use_nn"], cfg["show_non_nn"], cfg["density_factor"], cfg["g"], 1, 1)
    sc.set_dir_light((1, 1, 1), DIR_ZENITH, cfg["dir_dir"], DIR_AZIMUTH, cfg["dir_strength"])
    sc.set_point_light(cfg["pt_pos"], cfg["pt_strength"], cfg["pt_color"])
    sc.set_envmap(envmap(), None, None, cfg["env_direct"], cfg["env_hpm"])
    sc.set_mlp(S.mlp_init(1) if params is None else params, *lr)
    sc.set_mrhe(S.table_init(2, 16, T, 2) if tables is None else tables, res, mrhe_lr[0], mrhe_lr[1], 16, T, 2)
    return sc


def rand_queries(n, seed=0):
    g = np.random.Generator(np.random.PCG64(seed))
    pos = ((g.random((n, 3)) - 0.5) * SKY).astype(np.float32)
    d = g.standard_normal((n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return pos, d


def psnr(a, b, peak=None):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    mse = np.mean((a - b) ** 2)
    peak = max(np.abs(b).max(), 1e-12) if peak is None else peak
    return 99.0 if mse == 0 else 10.0 * np.log10(peak * peak / mse)
