"""GPU: k_mlp_stream (pipelined MLP-only tcgen05 kernel) -- agreement with the fused k_infer_tc and a timing sweep over the
number of tiles in flight per CTA (NRC_MLP_SLOTS) and the batch size.  Prints one JSON line per measurement."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import neuralradiancecaching_b200 as N   # noqa: E402
from neuralradiancecaching_b200 import scene as S   # noqa: E402

FLOP = 41344.0


def main():
    ctx = N.Context(0)
    ctx.set_precision(N._lib.MLP_BF16_TC)
    dev = torch.device("cuda", 0)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
    p = S.mlp_init(1); p[:20672] *= 10.0; p[20672:] = S.init_params(7, 323, 0.01)
    nrc.set_params(p)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(S.table_init(2))
    g = np.random.Generator(np.random.PCG64(5))
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    for n in (1000, 1 << 18, 1 << 21):
        pos = ((g.random((n, 3)) - 0.5) * np.array([62.5, 42.5, 76.5])).astype(np.float32)
        d = g.standard_normal((n, 3)); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
        dpos, dd = torch.from_numpy(pos).to(dev), torch.from_numpy(d).to(dev)
        ctx.set_infer_mode(False)
        fused = N.infer(nrc, m, dpos, dd); ctx.sync()
        tiles = N.encode_tiles(m, dpos, dd); ctx.sync()
        out = torch.empty((n, 3), dtype=torch.float32, device=dev)
        for slots in (5, 4, 3, 2, 1):
            os.environ["NRC_MLP_SLOTS"] = str(slots)
            out.zero_()
            N.mlp_forward_tiles(nrc, tiles, n, out=out); ctx.sync()
            diff = float((out - fused).abs().max()); ref = float(fused.abs().max())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            for _ in range(3):
                N.mlp_forward_tiles(nrc, tiles, n, out=out)
            with torch.cuda.stream(ext):
                e0.record()
            for _ in range(reps):
                N.mlp_forward_tiles(nrc, tiles, n, out=out)
            with torch.cuda.stream(ext):
                e1.record()
            ctx.sync()
            ms = e0.elapsed_time(e1) / reps
            tf = n * FLOP / (ms * 1e-3) / 1e12
            print(json.dumps({"n": n, "slots": slots, "max_abs_diff_vs_fused": diff, "max_abs_fused": ref, "bit_identical": bool(torch.equal(out, fused)),
                              "ms": ms, "TFLOP/s": tf, "frac_of_measured_bf16_burst": tf / peaks["bf16_tflops"],
                              "frac_of_measured_bf16_sustained": tf / peaks["bf16_tflops_sustained"],
                              "GB/s": n * 140.0 / (ms * 1e-3) / 1e9}), flush=True)
        # the two kernels of the split path and the fused kernel, per launch
        ctx.profile(True); ctx.profile_read(reset=True)
        os.environ["NRC_MLP_SLOTS"] = "5"
        for mode in (False, True):
            ctx.set_infer_mode(mode)
            for _ in range(10):
                N.infer(nrc, m, dpos, dd, out=out)
        ctx.sync()
        prof = ctx.profile_read(reset=True); ctx.profile(False)
        print(json.dumps({"n": n, "per_launch_ms": {k: v[0] / v[1] for k, v in prof.items()}}), flush=True)
    ctx.destroy()


if __name__ == "__main__":
    main()
