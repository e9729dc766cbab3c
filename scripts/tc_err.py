import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests")); sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, common as Cm
import neuralradiancecaching_b200 as N
ctx = N.Context(0); ctx.set_precision(N._lib.MLP_BF16_TC)
def rel(a,b): return float(np.max(np.abs(a.astype(np.float64)-b.astype(np.float64)))/max(float(np.max(np.abs(b))),1e-30))
for wscale in (0.01, 0.1):
  for n in (128, 3000, 1<<15):
    params = Cm.S.mlp_init(1); params[:20672] *= wscale/0.01; params[20672:] = Cm.S.init_params(7, 323, 0.01)
    tables = Cm.S.table_init()
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=params, tables=tables)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(params)
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(tables)
    pos, d = Cm.rand_queries(n, seed=21+n)
    target = (np.random.Generator(np.random.PCG64(4)).random((n,3))*3).astype(np.float32)
    sc.set_loss_scale(1.0/n, 1.0)
    op,_ = sc.train_accumulate(pos, d, target)
    gp = N.train_accumulate(nrc, m, pos, d, target, 1.0/n, want_pred=True)
    od = sc.get_mlp()[1]; gd = nrc.get_params()[1]
    e = ["%.3f" % rel(gd[l*4096:l*4096+(4096 if l<5 else 192)], od[l*4096:l*4096+(4096 if l<5 else 192)]) for l in range(6)]
    eb = ["%.3f" % rel(gd[20672+64*l:20672+64*l+(64 if l<5 else 3)], od[20672+64*l:20672+64*l+(64 if l<5 else 3)]) for l in range(6)]
    otd = sc.get_mrhe()[1][:32768]; gtd = m.get_tables()[1][:32768]
    print("w=%.2f n=%6d pred %.4f dW %s db %s tab %.3f loss %.4g vs %.4g" % (wscale, n, rel(gp, op), e, eb, rel(gtd, otd), nrc.GetStats().mseLoss, sc.mse_loss()))
    nrc.Destroy(); m.Destroy()
