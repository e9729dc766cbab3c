"""How much of k_render_walk's lane loss comes from paths of different length inside a warp?  Runs the device walk code on the host
(tests/host_emul) over the bench frame and compares, per warp, the lanes' walk steps and SDF evaluations with the warp maximum, for the
32x1 pixel-to-lane mapping the kernel uses and for 8x4 tiles.  No GPU needed.

  python scripts/walk_divergence.py [width height]
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from host_emul import emul as E  # noqa: E402


def main():
    W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
    inp = bench.scene_inputs(W, H)
    shape = E.set_volume(inp["density"])
    cam = inp["cam"]
    rnd = bench.lcg_randoms(2)[1]
    cfg = dict(cam=cam, random4=rnd, g=0.7, use_nn=1, density_factor=0.4)
    cost, per = E.render_walk_cost(cfg, W, H, shape)
    steps, sdf, camsdf, lit = [cost[..., k].astype(np.int64) for k in range(4)]
    print("pixels %d, in the box %.1f %%, mean walk steps %.2f (of those in the box %.2f), SDF evaluations per step %.2f, camera ray %.2f"
          % (W * H, 100.0 * (steps > 0).mean(), steps.mean(), steps[steps > 0].mean(), sdf.sum() / max(steps.sum(), 1), camsdf.mean()))

    def report(name, groups):      # groups: (nWarps, 32) index arrays into the flattened frame
        s, d = steps.reshape(-1)[groups], sdf.reshape(-1)[groups]
        live = s.max(axis=1) > 0
        q = per.reshape(-1, 32)[groups].astype(np.int64)          # (warps, lanes, steps)
        lock = q.max(axis=1).sum()                                 # what the warp executes: per step the slowest lane's march
        print("%-10s steps: lane sum / (32 x warp max) = %.3f | SDF evaluations, lanes / (32 x per-step warp max) = %.3f "
              "(if the paths were equally long: %.3f)"
              % (name, s.sum() / (32.0 * s.max(axis=1).sum()), q.sum() / (32.0 * lock), d.sum() / (32.0 * d.max(axis=1).sum())))

    idx = np.arange(W * H).reshape(H, W)
    report("32x1", idx.reshape(-1, 32))
    Ht, Wt = H - H % 4, W - W % 8
    t = idx[:Ht, :Wt].reshape(Ht // 4, 4, Wt // 8, 8).transpose(0, 2, 1, 3).reshape(-1, 32)
    report("8x4", t)
    # lanes sorted by cost: the bound any regrouping of whole paths could reach
    order = np.argsort(steps.reshape(-1), kind="stable")
    report("sorted", order[: (W * H) // 32 * 32].reshape(-1, 32))
    # distribution of walk steps
    # the two-launch walk (trace.cu: k_render_walk + k_render_walk_tail): head = steps [0, K) with the 32x1 mapping, tail = the paths
    # alive after K steps, compacted in warp order, steps [K, 32)
    q = per.reshape(-1, 32).astype(np.int64)
    st = steps.reshape(-1)
    one = q[idx.reshape(-1, 32)].max(axis=1).sum()
    for K in (2, 4, 6, 8, 10, 12, 16):
        head = q[idx.reshape(-1, 32)][:, :, :K].max(axis=1).sum()
        alive = np.nonzero((st == 32) | (st > K))[0]                   # a path that ends at step <= K never reaches the tail
        alive = alive[q[alive, K:].sum(axis=1) > 0] if K < 32 else alive[:0]
        pad = (-len(alive)) % 32
        a = np.concatenate([alive, np.full(pad, alive[-1])]).reshape(-1, 32)
        tq = q[a][:, :, K:]
        if pad:
            tq[-1, 32 - pad:, :] = 0
        tail = tq.max(axis=1).sum()
        print("split at %2d: head %5.1f %% + tail %5.1f %% = %5.1f %% of the single launch's warp-level SDF evaluations (%.1f %% of the paths reach the tail)"
              % (K, 100.0 * head / one, 100.0 * tail / one, 100.0 * (head + tail) / one, 100.0 * len(alive) / len(st)))
    hs = np.bincount(per.reshape(-1)[per.reshape(-1) > 0], minlength=40)
    print("SDF evaluations per step, histogram (1..):", " ".join("%.1f" % (100.0 * x / hs.sum()) for x in hs[1:40]))
    h = np.bincount(steps.reshape(-1), minlength=33)
    print("walk steps histogram (0..32):", " ".join("%.1f" % (100.0 * x / (W * H)) for x in h))


if __name__ == "__main__":
    main()
