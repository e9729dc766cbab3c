"""CPU emulation (numpy, fp64 accumulate) of the tensor-core train kernel's arithmetic against the fp32 oracle: which operand precision
does the hash-table gradient need?  Runs here without a GPU.  Result (round 2): the error comes from the FORWARD pass (bf16 rounding
flips ReLU gates that the whole backward pass reads), not from the last back-data GEMM; a three-part bf16 split of the forward
operands (six products per K step) brings the table gradient to the floor set by the bf16 backward pass (<= 0.75 %), a two-part
split leaves up to 6 %, plain bf16 12-30 %.  csrc/mlp_tc.cu (k_train_tc) implements the three-part split."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'tests')); sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
import numpy as np, common as Cm

def bf(x):
    x = np.asarray(x, np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return r.astype(np.uint32).view(np.float32)

def split(x):
    hi = bf(x); lo = bf(x - hi); return hi, lo

def run(n, wscale, mode, seed=21):
    params = Cm.S.mlp_init(1); params[:20672] *= wscale / 0.01; params[20672:] = Cm.S.init_params(7, 323, 0.01)
    tables = Cm.S.table_init()
    sc = Cm.oracle_scene(Cm.scene_cfg(), params=params, tables=tables)
    pos, d = Cm.rand_queries(n, seed=seed + n)
    target = (np.random.Generator(np.random.PCG64(4)).random((n, 3)) * 3).astype(np.float32)
    scale = 1.0 / n
    sc.set_loss_scale(scale, 1.0)
    sc.train_accumulate(pos, d, target)
    od = sc.get_mlp()[1]; otd = sc.get_mrhe()[1][:32768].astype(np.float64)
    feat, idx, lerp = sc.encode_pos(pos)
    x0 = np.concatenate([feat, sc.encode_dir(d)], 1).astype(np.float32)
    W = [params[l*4096:(l+1)*4096].reshape(64, 64) for l in range(5)] + [params[5*4096:5*4096+192].reshape(3, 64)]
    b = [params[20672+64*l:20672+64*l+64] for l in range(5)] + [params[20992:20995]]
    res = {}
    for name, cfg in mode.items():
        fw, bw, last, egate = cfg   # precision of forward operands, backward operands, last GEMM, E rounding
        def q(x, p):
            return bf(x) if p == 'bf16' else np.asarray(x, np.float32)
        def mm(a, w, p):   # a @ w.T with operand precision p, fp32 accumulate (emulated in float64 then cast)
            if p == 'split3':
                ah, r = bf(a), None
                am = bf(a - ah); al = bf(a - ah - am)
                wh = bf(w); wm = bf(w - wh); wl = bf(w - wh - wm)
                f = lambda x: x.astype(np.float64)
                return (f(ah) @ f(wh).T + f(ah) @ f(wm).T + f(am) @ f(wh).T + f(ah) @ f(wl).T + f(al) @ f(wh).T + f(am) @ f(wm).T).astype(np.float32)
            if p == 'split':  # hi*hi + hi*lo + lo*hi
                ah, al = split(a); wh, wl = split(w)
                return (ah.astype(np.float64) @ wh.T.astype(np.float64) + ah.astype(np.float64) @ wl.T.astype(np.float64) + al.astype(np.float64) @ wh.T.astype(np.float64)).astype(np.float32)
            return (q(a, p).astype(np.float64) @ q(w, p).T.astype(np.float64)).astype(np.float32)
        A = [q(x0, fw)]
        for l in range(5):
            A.append(q(np.maximum(mm(A[l], W[l], fw) + b[l], 0), fw))
        y = np.maximum(mm(A[5], W[5], fw) + b[5], 0)
        t = np.minimum(target, 1024.0)
        err = y - t
        e6 = np.where(A[5][:, :3] > 0, 2 * err, 0).astype(np.float32)
        E = q(e6, bw)
        dX = mm(E, W[5].T, bw)          # [n,64]
        for L in range(4, -1, -1):
            E = q(np.where(A[L] > 0, dX, 0), bw if L > 0 or True else bw)   # E_{L+1} gated by A_L
            if L == 0 and last != bw:
                E = np.where(A[0] > 0, dX, 0).astype(np.float32)            # keep E_1 unrounded for the precise last GEMM
            dX = mm(E, W[L].T, last if L == 0 else bw)
        g = dX  # [n,64]
        # table gradient
        lf = lerp.astype(np.float32)       # [n,16,3]
        tab = np.zeros(16384, np.float64)
        for lv in range(16):
            ng = -g[:, 2*lv].astype(np.float32)
            for c in range(8):
                xf = lf[:, lv, 0] if (c & 4) else (1.0 - lf[:, lv, 0])
                yf = lf[:, lv, 1] if (c & 2) else (1.0 - lf[:, lv, 1])
                zf = lf[:, lv, 2] if (c & 1) else (1.0 - lf[:, lv, 2])
                dl = ((ng * ((xf * yf) * zf)) * np.float32(scale)).astype(np.float64)
                np.add.at(tab, idx[:, lv*8 + c], dl)
        ref = otd[0::2]
        res[name] = np.max(np.abs(tab - ref)) / np.max(np.abs(ref))
    return res

modes = {
  'bf16 all': ('bf16', 'bf16', 'bf16', None),
  'bf16, last GEMM split': ('bf16', 'bf16', 'split', None),
  'split2 fwd': ('split', 'bf16', 'bf16', None),
  'split3 fwd': ('split3', 'bf16', 'bf16', None),
}
modes_unused = {
  'fp32 all': ('fp32', 'fp32', 'fp32', None),
  'bf16 all': ('bf16', 'bf16', 'bf16', None),
  'bf16, last GEMM split': ('bf16', 'bf16', 'split', None),
  'bf16 fwd, split bwd': ('bf16', 'split', 'split', None),
  'split fwd, bf16 bwd': ('split', 'bf16', 'bf16', None),
  'split fwd, bf16 bwd, last split': ('split', 'bf16', 'split', None),
  'split all': ('split', 'split', 'split', None),
}
for ws in (0.1, 0.01):
  for seed in (21, 5, 99):
    for n in (128, 3000, 1 << 15):
        r = run(n, ws, modes, seed)
        print('w=%.2f n=%6d' % (ws, n), {k: '%.4f' % v for k, v in r.items()})
