"""GPU: where does k_mlp_stream's time go?  Needs a build with NRC_EXTRA_NVCC_FLAGS=-DNRC_MLP_PROBES.  Each NRC_MLP_DBG value removes
one ingredient (bit 0 tcgen05.ld, bit 1 the bias/ReLU/convert arithmetic, bit 2 tcgen05.st, bit 3 the MMAs; 16 = both loads of
a row in flight; 116/216 = that with 4/3 slots); results of the probes are wrong by construction, only the time matters."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import neuralradiancecaching_b200 as N   # noqa: E402
from neuralradiancecaching_b200 import scene as S   # noqa: E402


def main():
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    ctx = N.Context(0)
    ctx.set_precision(N._lib.MLP_BF16_TC)
    dev = torch.device("cuda", 0)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(S.mlp_init(1))
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(S.table_init(2))
    g = np.random.Generator(np.random.PCG64(5))
    n = 1 << 21
    pos = ((g.random((n, 3)) - 0.5) * np.array([62.5, 42.5, 76.5])).astype(np.float32)
    d = g.standard_normal((n, 3)); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    dpos, dd = torch.from_numpy(pos).to(dev), torch.from_numpy(d).to(dev)
    tiles = N.encode_tiles(m, dpos, dd); ctx.sync()
    out = torch.empty((n, 3), dtype=torch.float32, device=dev)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    for dbg in sys.argv[1:] or ["0", "1", "2", "3", "4", "7", "8", "15", "16", "116", "216"]:
        os.environ["NRC_MLP_DBG"] = dbg
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(5):
            N.mlp_forward_tiles(nrc, tiles, n, out=out)
        reps = 200
        with torch.cuda.stream(ext):
            e0.record()
        for _ in range(reps):
            N.mlp_forward_tiles(nrc, tiles, n, out=out)
        clk = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
        with torch.cuda.stream(ext):
            e1.record()
        ctx.sync()
        ms = e0.elapsed_time(e1) / reps
        cyc = ms * 1e-3 * clk * 1e6 * 148 / (n / 128) / 6
        print(json.dumps({"dbg": int(dbg), "ms": ms, "TFLOP/s": n * 41344.0 / (ms * 1e-3) / 1e12, "sm_mhz": clk, "cycles_per_tile_layer": cyc}), flush=True)
    ctx.destroy()


if __name__ == "__main__":
    main()
