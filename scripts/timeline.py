"""Print the kernel timeline (both streams) of one frame: python scripts/timeline.py [bf16|fp32]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import neuralradiancecaching_b200 as N
from neuralradiancecaching_b200 import _lib as L

W, H = 1920, 1080
inp = bench.scene_inputs(W, H)
ctx = N.Context(0)
if len(sys.argv) > 1 and sys.argv[1] == "bf16":
    ctx.set_precision(L.MLP_BF16_TC)
vol = N.VolumeData(ctx, inp["density"]); vol.useNN = 1
env = N.HdrEnvMap(ctx, 1024, 512, inp["env"])
nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(inp["params"])
m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(inp["tables"])
r = N.NrcHpmRenderer(W, H, 100, 100, inp["cam"], vol, N.DirLight(-1.57, 0.0, (1, 1, 1), 1.0), N.PointLight((0, 0, 0), (1, 1, 1), 1.0), env, nrc, m)
rnd = bench.lcg_randoms(8)
for i in range(5):
    vol.Update(random4=rnd[i]); r.Render()
ctx.sync()
ctx.profile(True)
for i in range(2):
    vol.Update(random4=rnd[5 + i]); r.Render()
ctx.sync()
for name, s, e in ctx.profile_timeline():
    print("%-18s %8.3f -> %8.3f  (%.3f ms)" % (name, s, e, e - s))
