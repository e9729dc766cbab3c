"""GPU: a few launches of the split inference path's two kernels at one size (for ncu: -k regex:"k_mlp_stream|k_encode_tiles")."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import neuralradiancecaching_b200 as N   # noqa: E402
from neuralradiancecaching_b200 import scene as S   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 21
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    ctx = N.Context(0)
    ctx.set_precision(N._lib.MLP_BF16_TC)
    dev = torch.device("cuda", 0)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(S.mlp_init(1))
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(S.table_init(2))
    g = torch.Generator(device=dev); g.manual_seed(1)
    pos = ((torch.rand((n, 3), generator=g, device=dev) - 0.5) * torch.tensor([62.5, 42.5, 76.5], device=dev)).contiguous()
    d = torch.randn((n, 3), generator=g, device=dev); d = (d / d.norm(dim=1, keepdim=True)).contiguous()
    tiles = torch.empty(N.feature_tile_bytes(n), dtype=torch.uint8, device=dev)
    rgb = torch.empty((n, 3), dtype=torch.float32, device=dev)
    torch.cuda.synchronize(dev)
    for _ in range(reps):
        N.encode_tiles(m, pos, d, out=tiles)
        N.mlp_forward_tiles(nrc, tiles, n, out=rgb)
    ctx.sync()
    print("ok", n, float(rgb.sum()))
    ctx.destroy()


if __name__ == "__main__":
    main()
