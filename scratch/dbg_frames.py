import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, common as Cm
import neuralradiancecaching_b200 as N
from test_gpu_frame import build, _relmax
ctx = N.Context(0)
W, H, tw, th = 64, 48, 16, 16
cfg = Cm.scene_cfg(aspect=W / H)
sc = Cm.oracle_scene(cfg)
r, nrc, m, vol = build(N, ctx, cfg, W, H, tw, th)
rnd = [(0.37, 0.61, 0.13, 0.89), (0.11, 0.93, 0.5, 0.2), (0.71, 0.29, 0.4, 0.6), (0.3,0.3,0.3,0.3)]
for f in range(4):
    sc.set_volume_params(rnd[f], 1, 0, 0.4, 0.7, 1, 1)
    sc.reset_stats()
    # oracle train part only, to look at gradients
    sc.set_train_size(tw, th)
    orec, _ = sc.train_trace(tw, th)
    sc.train_accumulate(orec[:, 0:3], orec[:, 3:6], orec[:, 6:9])
    og = sc.get_mlp()[1].copy(); otg = sc.get_mrhe()[1].copy()
    vol.random = np.asarray(rnd[f], np.float32); nrc.ResetStats()
    r.FrameTrain(); ctx.sync()
    gg = nrc.get_params()[1]; gtg = m.get_tables()[1]
    rec, _ = r.read_train_probes()
    print(f, "rec bit-eq", np.array_equal(rec[:, :6].view(np.uint32), orec[:, :6].view(np.uint32)), "target rel", _relmax(rec[:,6:], orec[:,6:]),
          "grad rel", _relmax(gg, og), "grad absmax", np.abs(og).max(), "n clipped", (np.abs(og) > 1000).sum(), "tgrad rel", _relmax(gtg[:32768], otg[:32768]), np.abs(otg).max())
    N.optimizer_step(nrc, m); sc.nrc_step(); sc.mrhe_step()
    print("   params rel", _relmax(nrc.get_params()[0], sc.get_mlp()[0]), "tables rel", _relmax(m.get_tables()[0], sc.get_mrhe()[0]), "loss", nrc.GetStats().mseLoss, sc.mse_loss())
    r.FrameRender(); ctx.sync()
    frame = r.read_frame()
    oframe, _, oq = sc.render(W, H)
    cnt, q = r.read_render_probes()
    print("   psnr", Cm.psnr(frame, oframe), "frame max", oframe[...,:3].max(), "relmax", _relmax(frame, oframe))
    valid = oq[..., 7] > 0
    qp, qd = oq[..., :3][valid], oq[..., 3:6][valid]
    oi = sc.infer(qp, qd); gi = N.infer(nrc, m, qp, qd)
    err = np.abs(frame - oframe).max(-1)
    print("   infer out max", np.abs(oi).max(), "infer abs err", np.abs(oi - gi).max(), "frame err on valid px", err[valid].max(), "on others", err[~valid].max() if (~valid).any() else 0)
    k = np.argmax(np.abs(oi - gi).max(1)); print("   worst query", qp[k], qd[k], oi[k], gi[k])
    f, _, _ = sc.encode_pos(qp[k:k+1]); print("   feat absmax", np.abs(f).max())
