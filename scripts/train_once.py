"""GPU: a few nrc_train_accumulate launches on device-resident 2^18-sample batches (BASELINE config 2), for ncu (-k regex:k_train_tc)
and for timing the phases of the train kernel (NRC_TRAIN_PHASES knob, if the build has it)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import neuralradiancecaching_b200 as N   # noqa: E402
from neuralradiancecaching_b200 import scene as S   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
    ctx = N.Context(0)
    if prec == "bf16":
        ctx.set_precision(N._lib.MLP_BF16_TC)
    dev = torch.device("cuda", 0)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(S.mlp_init(1))
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(S.table_init(2))
    g = torch.Generator(device=dev); g.manual_seed(1)
    pos = ((torch.rand((n, 3), generator=g, device=dev) - 0.5) * torch.tensor([62.5, 42.5, 76.5], device=dev)).contiguous()
    d = torch.randn((n, 3), generator=g, device=dev); d = (d / d.norm(dim=1, keepdim=True)).contiguous()
    t = torch.rand((n, 3), generator=g, device=dev).contiguous()
    torch.cuda.synchronize(dev)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    N.train_accumulate(nrc, m, pos, d, t, 1.0 / n)
    N.optimizer_step(nrc, m)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(ext):
        e0.record()
    for _ in range(reps):
        N.train_accumulate(nrc, m, pos, d, t, 1.0 / n)
    with torch.cuda.stream(ext):
        e1.record()
    ctx.sync()
    print("train_accumulate n=%d %s: %.4f ms per launch" % (n, prec, e0.elapsed_time(e1) / reps))


if __name__ == "__main__":
    main()
