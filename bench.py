"""bench.py -- headline benchmark of the NRC hot path (BASELINE.json: NRC queries/s and ms/frame, cloud 1080p).

A step = one frame of NrcHpmRenderer::Render (train trace -> encode+MLP fwd/bwd -> nrc-step -> mrhe-step -> render
trace -> batched NRC inference -> composite) on the synthetic cloud_sixteenth stand-in (the real blob is missing from
the reference, .MISSING_LARGE_BLOBS).  At N GPUs the frame holds N x 1080p pixels (16:9 kept, both sides scaled by
sqrt(N)), split into interleaved row sets, and the train grid holds N x 10 000 samples, sharded; gradients are
all-reduced over NCCL (weak scaling).  value = (inference queries + training samples) of all ranks / max-over-ranks
device time.  `e2e` is the same through the host-facing call sequence with the frame read back to pinned host memory
every step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload frame|fit]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "nrc_queries_per_s"
UNIT = "queries/s"
BASE_W, BASE_H, BASE_TW, BASE_TH = 1920, 1080, 100, 100


def frame_shape(n_gpus):
    from neuralradiancecaching_b200.dist import frame_shape as fs
    return fs(n_gpus, (BASE_W, BASE_H, BASE_TW, BASE_TH))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
        except Exception:
            pass
    return 6650.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx = max(mx, float(s[1]))
                for n, v in zip(names, s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- scene shared by both arms -------------------------------------------------------------------------------------
def scene_inputs(width, height, grid_scale=1, table_size=16384):
    from neuralradiancecaching_b200 import scene as S
    from neuralradiancecaching_b200.objects import Camera, vec_from_angles
    dens = S.cloud_density()                      # 125 x 85 x 153, deterministic (SHA-256 in BASELINE.md)
    if grid_scale > 1:                            # config 5: the "full" cloud = the sixteenth-size grid refined per axis
        dens = S.refined_density(grid_scale)
    env_path = os.environ.get("NRC_ENVMAP")       # e.g. the reference's data/image/photostudio.hdr (1024 x 512 RGBE); not present on the GPU box
    if env_path:
        from neuralradiancecaching_b200 import loaders as Ld
        env, ew, eh = Ld.ReadFileHdr4f(env_path)
        if (ew, eh) != (1024, 512):
            raise SystemExit("bench.py: NRC_ENVMAP must be a 1024x512 map (the shape of data/image/photostudio.hdr), got %dx%d" % (ew, eh))
    else:
        env = S.sky_envmap(1024, 512)             # the shape of data/image/photostudio.hdr
    cam = Camera((0.0, 0.0, -64.0), (0.0, 0.0, 1.0), (0.0, 1.0, 0.0), width / float(height), np.radians(60.0), 0.1, 100.0)
    return dict(density=dens, env=env, cam=cam, dir_dir=vec_from_angles(-1.57, 0.0), params=S.mlp_init(1), tables=S.table_init(2, 16, table_size, 2))


def lcg_randoms(n, seed=0x2545F491):
    """the per-frame `random` vec4 (VolumeData::Update, src/VolumeData.cpp:105-109) from a seeded LCG"""
    out, s = [], seed
    for _ in range(n):
        r = []
        for _ in range(4):
            s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
            r.append((s >> 8) / float(1 << 24))
        out.append(np.asarray(r, np.float32))
    return out


# ---- the B200 arm ------------------------------------------------------------------------------------------------------
class Workload:
    """One BASELINE configuration set up on this rank's GPU: scene objects, renderer, this rank's partition, the gradient exchange."""

    def __init__(self, args, ctx, rank, world, local, W, H, TW, TH, grid_scale=1, table_size=16384, exchange=None):
        import torch
        import neuralradiancecaching_b200 as N
        from neuralradiancecaching_b200 import _lib as L
        from neuralradiancecaching_b200.dist import GradientExchange, FusedExchange
        self.N, self.L, self.ctx, self.rank, self.world, self.local = N, L, ctx, rank, world, local
        self.W, self.H, self.TW, self.TH, self.grid_scale, self.table_size = W, H, TW, TH, grid_scale, table_size
        self.inp = inp = scene_inputs(W, H, grid_scale, table_size)
        self.vol = N.VolumeData(ctx, inp["density"])
        self.vol.useNN = 1
        self.env = N.HdrEnvMap(ctx, 1024, 512, inp["env"])
        self.dl = N.DirLight(-1.57, 0.0, (1, 1, 1), 1.0)      # strengths 1.0: the reference default 0 makes the target == 0 (SURVEY finding 9)
        self.pl = N.PointLight((0, 0, 0), (1, 1, 1), 1.0)
        self.nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0)
        self.nrc.set_params(inp["params"])
        self.mrhe = N.MRHE(ctx, 0.01, 0.0, 16, table_size, 2, seed=0)
        self.mrhe.set_tables(inp["tables"])
        self.r = N.NrcHpmRenderer(W, H, TW, TH, inp["cam"], self.vol, self.dl, self.pl, self.env, self.nrc, self.mrhe)
        if world > 1:
            self.r.SetPartitionStrided(rank, world, rank, world)
        # Render vertex pool, vertices per path.  Left to the library unless NRC_BENCH_POOL=n is set: it takes the worst case (32 per path:
        # cannot overflow, so no frame waits on the host for its overflow flag; 3 x 8.5 GB of HBM per 1080p rank, which is what 180 GB are
        # for) when that is at most a sixth of the GPU's memory, else 4 per path with the deferred check (csrc/renderer.cu).
        self.pool_per_path = int(os.environ.get("NRC_BENCH_POOL", "0"))
        self.pool_from_env = "NRC_BENCH_POOL" in os.environ
        if self.pool_from_env and self.pool_per_path > 0:
            self.r.SetPoolCapacity(-(-H // world) * W * self.pool_per_path)      # the rank with the most rows
        elif self.pool_from_env:
            self.r.SetPoolCapacity(-(-H // world) * W * 4)                      # NRC_BENCH_POOL=0: force the small pool
        dev = torch.device("cuda", local)
        self.exchange = GradientExchange(ctx, self.nrc, self.mrhe, dev)
        self.ext = self.exchange.stream
        kind = exchange or args.exchange
        self.exchange_kind = kind if world > 1 else "none"
        self.fused = FusedExchange(self.nrc, self.mrhe, rank, world, dev) if (world > 1 and kind == "fused") else None

    def frame(self, random4):
        self.vol.Update(random4=random4)
        self.nrc.ResetStats()
        if self.world == 1:
            self.r.Render()
        else:
            self.r.FrameTrain()
            if self.fused is not None:
                self.fused()                 # one kernel: peer-memory sum over NVLink + nrc-step + mrhe-step (csrc/dp.cu)
            else:
                self.exchange()              # NCCL all-reduce of the MLP gradient arena (+loss) and the delta-table head
                self.N.optimizer_step(self.nrc, self.mrhe)
            self.r.FrameRender()

    def barrier(self):
        import torch
        import torch.distributed as dist
        self.ctx.sync()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(self, K, warm, rnd_iter, profile=False):
        """`warm` untimed frames, then K frames between two events on the main stream, barrier + synchronize on both sides; returns
        (ms of the K frames on this rank, this rank's counters over them, per-kernel profile or None, kernel launches)."""
        import torch
        ctx = self.ctx
        for _ in range(warm):
            self.frame(next(rnd_iter))
        self.barrier()
        self.r.totals(reset=True)
        launches0 = ctx.launch_count()
        if profile:
            ctx.profile(True)
            ctx.profile_read(reset=True)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self.ext):
            ev0.record()
        for _ in range(K):
            self.frame(next(rnd_iter))
        with torch.cuda.stream(self.ext):
            ev1.record()
        self.barrier()
        ms = ev0.elapsed_time(ev1)
        prof = None
        if profile:
            prof = ctx.profile_read(reset=True)
            ctx.profile(False)
        return ms, self.r.totals(reset=True), prof, ctx.launch_count() - launches0

    def check_exchange(self):
        if self.fused is not None and self.fused.timed_out():
            raise SystemExit("bench.py: a handshake of the fused gradient exchange timed out; the replicas may have diverged -- no result")

    def destroy(self):
        self.ctx.sync()
        if self.fused is not None:
            self.fused.close()
        for o in (self.r, self.nrc, self.mrhe, self.vol, self.env):
            o.Destroy()


def side_measure(args, ctx, rank, world, local, W, H, TW, TH, grid_scale, table_size, scaling, steps=20, warm=6, label=""):
    """A short measurement of another BASELINE configuration after the headline loop (own objects, same context): ms/frame, queries/s
    over all ranks (max-over-ranks device time), the slowest kernels of this rank, and its own oracle parity flags."""
    import torch
    import torch.distributed as dist
    wl = Workload(args, ctx, rank, world, local, W, H, TW, TH, grid_scale, table_size)
    rnd = iter(lcg_randoms(steps + warm + 4, seed=0x1F123BB5))
    ms, tot, prof, launches = wl.timed(steps, warm, rnd, profile=True)
    wl.check_exchange()
    dev = torch.device("cuda", local)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    c = torch.tensor([tot["infer_queries"] + tot["train_samples"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c)
    out = {"workload": label, "frame": [W, H], "train_grid": [TW, TH], "hash_table": [16, table_size, 2], "grid_scale": grid_scale,
           "n_gpus": world, "scaling": scaling, "steps": steps, "warmup": warm, "ms_per_step": float(t[0]) / steps,
           "queries_per_s": float(c[0]) / (float(t[0]) * 1e-3), "exchange": wl.exchange_kind,
           "kernels_ms_per_step": {k: ms_ / steps for k, (ms_, _) in sorted(prof.items(), key=lambda kv: -kv[1][0])[:8]}}
    if not args.no_parity:
        try:
            v = {"r": wl.r, "nrc": wl.nrc, "mrhe": wl.mrhe, "vol": wl.vol, "ctx": ctx, "inp": wl.inp, "world": world, "rank": rank, "local": local,
                 "W": W, "H": H, "TW": TW, "TH": TH, "table_size": table_size}
            out["parity"] = parity_check(v, args)
        except Exception as e:
            out["parity"] = {"error": repr(e)}
    wl.destroy()
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    import neuralradiancecaching_b200 as N
    from neuralradiancecaching_b200 import _lib as L

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_gpus = world
    W, H, TW, TH = frame_shape(n_gpus)
    scaling, grid_scale, table_size = "weak", 1, 16384
    if args.config == "3":                        # 1280x720 on one GPU
        W, H, TW, TH = 1280, 720, 100, 100
    elif args.config == "4":                      # one 4K frame split over the ranks, the 10 000 samples sharded: strong scaling
        W, H, TW, TH, scaling = 3840, 2160, 100, 100, "strong"
    elif args.config == "5":                      # refined cloud (500x340x612), 2^22-entry hash tables, weak-scaling frame
        grid_scale, table_size = 4, 1 << 22
    if args.width:
        W, H = args.width, args.height

    ctx = N.Context(local)
    if args.precision == "bf16":
        ctx.set_precision(L.MLP_BF16_TC)
    wl = Workload(args, ctx, rank, world, local, W, H, TW, TH, grid_scale, table_size)
    inp, vol, env, nrc, mrhe, r = wl.inp, wl.vol, wl.env, wl.nrc, wl.mrhe, wl.r
    exchange, ext, fused, pool_per_path = wl.exchange, wl.ext, wl.fused, wl.pool_per_path
    args.exchange = wl.exchange_kind if world > 1 else args.exchange
    frame, barrier = wl.frame, wl.barrier

    K, Wm = args.steps, args.warmup
    REPS = max(1, args.repeats)
    rnd = lcg_randoms((2 * REPS + 2) * (K + Wm) + 60)
    ri = iter(rnd)
    dev = torch.device("cuda", local)

    def median_repeat(ms_list):
        """index of the median repeat by the max-over-ranks time (the same on every rank) and the reduced times"""
        t = torch.tensor(ms_list, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        order = sorted(range(len(ms_list)), key=lambda i: float(t[i]))
        return order[len(order) // 2], [float(x) for x in t]

    # ---- device-resident timing (value): REPS repeats of exactly K frames each, the median repeat is reported ----
    sampler = ClockSampler(local)      # rank 0 only: N concurrent nvidia-smi pollers perturb the ranks they poll
    reps = []
    for rep in range(REPS):
        if rep == 0:
            for _ in range(Wm):
                frame(next(ri))
            if rank == 0:
                sampler.start()
        reps.append(wl.timed(K, 0, ri, profile=True))
    sampler.stop_flag = True
    wl.check_exchange()
    mid, rep_ms = median_repeat([x[0] for x in reps])
    ms_total, tot, prof, launches = reps[mid]
    mse = nrc.GetStats().mseLoss

    # ---- end to end: the host-facing call sequence; per step the scene/camera/RNG structs go host->device (kernel
    # parameters) and this rank's rows of the BGRA8 frame plus the loss come back to pinned host memory ----
    # Like the reference's main loop (frame k is presented while frame k+1 is recorded, src/main.cpp:197-230) the read-back
    # is asynchronous and buffered: step k queues frame k and its copy, then takes delivery of frame k-D (D = NRC_E2E_DEPTH, default 2:
    # the host stays two frames ahead of the frame it consumes, like the device-timed loop, whose frame graph overlaps the next
    # frame's tracing with this frame's lighting and inference; D = 1 is the plain double buffer).  Every frame and every loss value
    # crosses PCIe inside the timed region; the last ones are awaited before the stop event.
    d2h_bytes = r.partition_bytes(L.FRAME_BGRA8)
    depth = min(max(int(os.environ.get("NRC_E2E_DEPTH", "2")), 1), 3)      # the renderer keeps 4 read-back tickets
    host_frames = [torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory() for _ in range(depth + 1)]
    host_loss = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(depth + 1)]
    hfs = [h.numpy() for h in host_frames]
    hls = [h.numpy() for h in host_loss]
    sync_e2e = bool(os.environ.get("NRC_E2E_SYNC"))

    def e2e_steps(count):
        pending = []
        for k in range(count):
            frame(next(ri))
            if sync_e2e:
                r.read_frame_partition(hfs[0], L.FRAME_BGRA8)    # synchronous D2H of this rank's rows
                _ = nrc.GetStats().mseLoss                        # D2H of the loss (src/main.cpp:220-224)
                continue
            nrc.get_stats_async(hls[k % (depth + 1)])
            pending.append(r.read_frame_partition_async(hfs[k % (depth + 1)], L.FRAME_BGRA8))
            if len(pending) > depth:
                r.readback_wait(pending.pop(0))                   # frame k-D and its loss are in host memory now
        for t in pending:
            r.readback_wait(t)

    e2e_steps(3)
    e2e_reps = []
    for rep in range(REPS):
        barrier()
        r.totals(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            e0.record()
        e2e_steps(K)
        ctx.sync()
        with torch.cuda.stream(ext):
            e1.record()
        barrier()
        e2e_reps.append((max(e0.elapsed_time(e1), 0.0), r.totals(reset=True)))
    wl.check_exchange()
    emid, e2e_rep_ms = median_repeat([x[0] for x in e2e_reps])
    e2e_ms, tot_e2e = e2e_reps[emid]
    if os.environ.get("NRC_BENCH_TIMELINE") and rank == 0:      # kernel timeline of two more frames (both streams), to stderr
        ctx.profile(True)
    if os.environ.get("NRC_BENCH_TIMELINE"):
        for _ in range(2):
            frame(next(ri))
        barrier()
        if rank == 0:
            for name, s0, s1 in ctx.profile_timeline():
                sys.stderr.write("TIMELINE %-18s %8.3f -> %8.3f  (%.3f ms)\n" % (name, s0, s1, s1 - s0))
            ctx.profile(False)
        r.totals(reset=True)
        if world > 1:      # the exchange alone, ranks aligned: pure collective cost
            barrier()
            x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(ext):
                x0.record()
            for _ in range(50):
                exchange()
            with torch.cuda.stream(ext):
                x1.record()
            barrier()
            if rank == 0:
                sys.stderr.write("TIMELINE exchange alone: %.1f us per frame (2 all-reduces)\n" % (x0.elapsed_time(x1) / 50 * 1e3))
            nrc.set_state(delta=np.zeros(20995, np.float32)); mrhe.set_tables(delta=np.zeros(mrhe.count, np.float32))
    # ---- the same frames with the fp32 CUDA-core MLP (the 1e-3 parity path), reported beside the headline ----
    fp32_ms = None
    if args.precision == "bf16" and not args.no_fp32_line:
        ctx.set_precision(L.MLP_FP32)
        for _ in range(3):
            frame(next(ri))
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            f0.record()
        kf = min(K, 10)
        for _ in range(kf):
            frame(next(ri))
        with torch.cuda.stream(ext):
            f1.record()
        barrier()
        fp32_ms = f0.elapsed_time(f1) / kf
        r.totals(reset=True)
        ctx.set_precision(L.MLP_BF16_TC)
    return finish_b200(args, locals())


# ---- BASELINE config 2: MRHE + 64-wide MLP fit to a synthetic 3-D radiance field, 2^18-query batches ------------------------
def synthetic_field(pos, d):
    """closed-form target: three Gaussian blobs in the volume times a forward-scattering lobe around a fixed sun direction"""
    centers = np.array([[-10.0, 2.0, 5.0], [8.0, -4.0, -12.0], [0.0, 6.0, 20.0]], np.float32)
    sig = np.array([9.0, 7.0, 11.0], np.float32)
    col = np.array([[1.0, 0.8, 0.6], [0.5, 0.9, 1.0], [0.9, 0.9, 0.9]], np.float32)
    sun = np.array([0.3, 0.8, -0.52], np.float32); sun /= np.linalg.norm(sun)
    out = np.zeros((pos.shape[0], 3), np.float32)
    for c, s_, k in zip(centers, sig, col):
        w = np.exp(-np.sum((pos - c) ** 2, axis=1) / (2.0 * s_ * s_)).astype(np.float32)
        out += w[:, None] * k[None, :]
    g = 0.7
    cos = d @ sun
    lobe = 0.5 * (1 - g * g) / np.power(1 + g * g - 2 * g * cos, 1.5)
    return (out * lobe[:, None]).astype(np.float32)


def run_fit(ctx, N, L, torch, steps=10, n=1 << 18):
    """steps/s and queries/s of train_accumulate + optimizer_step and of infer on device-resident 2^18 batches, plus the
    same through host buffers; per-kernel TFLOP/s from the library's events."""
    from neuralradiancecaching_b200 import scene as S
    dev = torch.device("cuda", ctx.device)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(S.mlp_init(1))
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(S.table_init(2))
    g = np.random.Generator(np.random.PCG64(0x5EED))
    batches = []
    for b in range(4):
        pos = ((g.random((n, 3)) - 0.5) * np.array([62.5, 42.5, 76.5])).astype(np.float32)
        d = g.standard_normal((n, 3)); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
        batches.append((pos, d, synthetic_field(pos, d)))
    devb = [tuple(torch.from_numpy(a).to(dev) for a in b) for b in batches]
    out = torch.empty((n, 3), dtype=torch.float32, device=dev)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def step(b, host=False):
        pos, d, t = (batches if host else devb)[b % 4]
        N.train_accumulate(nrc, m, pos, d, t, 1.0 / n)
        N.optimizer_step(nrc, m)

    for i in range(3):
        step(i)
    ctx.sync()
    losses = []
    ctx.profile(True); ctx.profile_read(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(ext):
        e0.record()
    for i in range(steps):
        step(i)
    with torch.cuda.stream(ext):
        e1.record()
    ctx.sync()
    train_ms = e0.elapsed_time(e1) / steps
    with torch.cuda.stream(ext):
        e0.record()
    for i in range(steps):
        N.infer(nrc, m, devb[i % 4][0], devb[i % 4][1], out=out)
    with torch.cuda.stream(ext):
        e1.record()
    ctx.sync()
    infer_ms = e0.elapsed_time(e1) / steps
    prof = ctx.profile_read(reset=True); ctx.profile(False)
    t0 = time.perf_counter()
    for i in range(steps):
        nrc.ResetStats()
        step(i, host=True)
        losses.append(nrc.GetStats().mseLoss)
    ctx.sync()
    host_ms = (time.perf_counter() - t0) * 1e3 / steps
    kern = {}
    for k, (ms, cnt) in prof.items():
        e = {"ms_per_launch": ms / cnt}
        if k in ("k_train_tc", "k_train_fp32"):
            e["TFLOP/s"] = n * FLOP_TRAIN / (ms / cnt * 1e-3) / 1e12
        if k in ("k_infer_tc", "k_infer_fp32"):
            e["TFLOP/s"] = n * FLOP_INFER / (ms / cnt * 1e-3) / 1e12
        if k == "k_mrhe_step":
            e["GB/s"] = 16.0 * m.count / (ms / cnt * 1e-3) / 1e9
        kern[k] = e
    nrc.Destroy(); m.Destroy()
    return {"workload": "MRHE(16x2^14x2)+MLP fit of a synthetic radiance field, %d-query batches" % n, "train_step_ms": train_ms,
            "train_queries_per_s": n / (train_ms * 1e-3), "infer_ms": infer_ms, "infer_queries_per_s": n / (infer_ms * 1e-3),
            "train_step_ms_host_buffers": host_ms, "h2d_bytes_per_step": n * 36, "loss_first_last": [losses[0], losses[-1]], "kernels": kern}


def run_mlp_stream(ctx, N, torch, sizes=(1 << 19, 1 << 21, 1 << 23), reps=20):
    """The split inference path's two kernels alone, inputs resident in HBM and larger than L2: k_encode_tiles (gather kernel,
    bounded by L2/HBM gathers: 1 036 algorithmic bytes per query) and k_mlp_stream (the six layers on tcgen05, bounded by the
    tensor pipe: 41 344 FLOP per query).  Durations are CUDA events on the launching stream around `reps` back-to-back launches."""
    from neuralradiancecaching_b200 import scene as S
    dev = torch.device("cuda", ctx.device)
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pk = {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1590.0, "hbm_gbs": 6650.0, "fallback": True}
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=0); nrc.set_params(S.mlp_init(1))
    m = N.MRHE(ctx, 0.01, 0.0, seed=0); m.set_tables(S.table_init(2))
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    g = torch.Generator(device=dev); g.manual_seed(0x5EED)
    out = {"peaks": {"bf16_tflops_burst": pk.get("bf16_tflops"), "bf16_tflops_sustained": pk.get("bf16_tflops_sustained"), "hbm_gbs": pk.get("hbm_gbs")},
           "sizes": []}

    def timed(fn):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            fn()
        with torch.cuda.stream(ext):
            e0.record()
        for _ in range(reps):
            fn()
        with torch.cuda.stream(ext):
            e1.record()
        ctx.sync()
        return e0.elapsed_time(e1) / reps

    for n in sizes:
        box = torch.tensor([62.5, 42.5, 76.5], device=dev)
        pos = ((torch.rand((n, 3), generator=g, device=dev) - 0.5) * box).contiguous()
        d = torch.randn((n, 3), generator=g, device=dev); d = (d / d.norm(dim=1, keepdim=True)).contiguous()
        tiles = torch.empty(N.feature_tile_bytes(n), dtype=torch.uint8, device=dev)
        rgb = torch.empty((n, 3), dtype=torch.float32, device=dev)
        torch.cuda.synchronize(dev)
        enc_ms = timed(lambda: N.encode_tiles(m, pos, d, out=tiles))
        mlp_ms = timed(lambda: N.mlp_forward_tiles(nrc, tiles, n, out=rgb))
        tf = n * FLOP_INFER / (mlp_ms * 1e-3) / 1e12
        out["sizes"].append({
            "queries": n, "feature_tile_MB": tiles.numel() / 1e6,
            "k_mlp_stream": {"ms": mlp_ms, "TFLOP/s": tf, "frac_of_bf16_burst": tf / pk["bf16_tflops"] if pk.get("bf16_tflops") else None,
                             "frac_of_bf16_sustained": tf / pk["bf16_tflops_sustained"] if pk.get("bf16_tflops_sustained") else None,
                             "GB/s": n * 140.0 / (mlp_ms * 1e-3) / 1e9, "queries_per_s": n / (mlp_ms * 1e-3)},
            "k_encode_tiles": {"ms": enc_ms, "GB/s_algorithmic": n * 1036.0 / (enc_ms * 1e-3) / 1e9, "queries_per_s": n / (enc_ms * 1e-3),
                               "note": "16x8 float2 table gathers + 24 B of pos/dir in, 128 B of bf16 features out per query; the 2 MiB table is L2-resident"}})
        del pos, d, tiles, rgb
    nrc.Destroy(); m.Destroy()
    return out


# algorithmic work per unit (DESIGN.md "Kernels")
TAP_BYTES = 8.0              # one density tap = one 2x2x2 UNORM8 brick
FLOP_INFER = 41344.0         # 2 * 20 672 MAC per query
FLOP_TRAIN = 124032.0        # forward + dW + dX per sample


# ncu --set full capture of the default N=1 workload: DRAM bytes (read + write) and executed warp instructions per launch of every
# frame kernel, read from the committed summary profiles/ncu_frame.json (written by scripts/ncu_summary.py --json from the capture named
# in its "source"); only quoted when the run IS that workload.
def vertex_pool_note(v):
    """What the render vertex pool of the timed frames could hold, per path of this rank, and who chose it."""
    my_rows = len(range(v["rank"], v["H"], v["world"]))
    per_path = v["r"].GetPoolCapacity() / max(my_rows * v["W"], 1)
    who = "NRC_BENCH_POOL" if "NRC_BENCH_POOL" in os.environ else "the library's choice"
    return "%.0f vertices per path (%s)%s" % (per_path, who, ": worst case, cannot overflow, no host wait per frame" if per_path >= 32 else
                                             ": a frame's overflow flag is read on the host before the next one is traced")


def load_ncu_frame():
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_frame.json")))
        return {k: (float(x[0]), float(x[1])) for k, x in d["kernels"].items()}, d.get("source", "profiles/ncu_frame.json")
    except Exception:
        return {}, None


NCU_FRAME, NCU_SOURCE = load_ncu_frame()


def kernel_work(name, tot, pixels, table_floats):
    """(bound, algorithmic units per timed region, unit) of one kernel, from the run's exact counters"""
    if name == "k_light<render>":
        return "hbm", tot["render_taps"] * TAP_BYTES + tot["render_vertices"] * 128.0, "GB/s"
    if name == "k_light<train>":
        return "hbm", tot["train_taps"] * TAP_BYTES + tot["train_vertices"] * 64.0, "GB/s"
    if name in ("k_infer_fp32", "k_infer_tc", "k_mlp_stream"):
        return "tensor", tot["infer_queries"] * FLOP_INFER, "TFLOP/s"
    if name == "k_encode_tiles":
        return "hbm", tot["infer_queries"] * (1036.0 + 128.0), "GB/s"      # 1 164 B/query stand-alone encode
    if name in ("k_train_fp32", "k_train_tc"):
        return "tensor", tot["train_samples"] * FLOP_TRAIN, "TFLOP/s"
    if name == "k_render_walk":      # head launch: background + path outputs per pixel, the vertices it emits; the spill is not algorithmic
        return "hbm", pixels * 68.0 + tot["render_vertices"] * 112.0, "GB/s"
    if name == "k_train_walk":
        return "hbm", tot["train_samples"] * 72.0 + tot["train_vertices"] * 48.0, "GB/s"
    if name == "k_mrhe_step":
        return "hbm", None, "GB/s"      # filled by the caller: 16 B/param/launch
    return "hbm", None, "GB/s"


def parity_check(v, args):
    """Oracle parity of the very workload the line times: the bench scene at its initial weights, one frame with a
    fixed `random` vec4, rendered once more on the GPU OUTSIDE the timed region and compared with the CPU oracle -- control flow bit for
    bit (RNG draws, lit vertices, walk steps, emitted queries, query / record positions and directions), the frame by PSNR, the
    gradients by max-norm.  N = 1: every train invocation and every `row_step`-th frame row; N > 1: every rank checks a sample of its
    own rows and of its own train invocations, and all ranks compare a hash of their parameters and tables (replicas_identical)."""
    import hashlib
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    r, nrc, mrhe, vol, ctx, inp = v["r"], v["nrc"], v["mrhe"], v["vol"], v["ctx"], v["inp"]
    world, rank, W, H, TW, TH = v["world"], v["rank"], v["W"], v["H"], v["TW"], v["TH"]
    out = {}
    # ---- replicas after the timed frames ----
    ctx.sync()
    if world > 1:
        hsh = hashlib.sha256(nrc.get_params()[0].tobytes() + mrhe.get_tables()[0].tobytes()).digest()[:8]
        mine = torch.tensor(list(hsh), dtype=torch.uint8, device=torch.device("cuda", v["local"]))
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        out["replicas_identical"] = all(bool((p == parts[0]).all()) for p in parts)
    # ---- the parity frame: initial weights, fixed seed ----
    rnd = np.asarray([0.37, 0.61, 0.13, 0.89], np.float32)
    nrc.set_params(inp["params"])
    nrc.set_state(delta=np.zeros(20995, np.float32), momentum=np.zeros(20995, np.float32))
    mrhe.set_tables(inp["tables"], delta=np.zeros(mrhe.count, np.float32))
    vol.Update(random4=rnd)
    nrc.ResetStats()
    r.FrameTrain()
    ctx.sync()
    rec, tcnt = r.read_train_probes()
    g_mlp = nrc.get_params()[1].copy()
    g_tab = mrhe.get_tables()[1][:2 * v["table_size"]].copy()
    if world == 1:
        import neuralradiancecaching_b200 as N
        N.optimizer_step(nrc, mrhe)
    else:      # the parity frame is per rank: no exchange; render with the initial weights
        nrc.set_state(delta=np.zeros(20995, np.float32)); mrhe.set_tables(delta=np.zeros(mrhe.count, np.float32))
    r.FrameRender()
    ctx.sync()
    cnt, qry = r.read_render_probes()
    frame = r.read_frame()
    # ---- the oracle on the same scene ----
    threads = max(1, (len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)) // world)
    O.set_num_threads(threads)
    sc = O.OracleScene()
    cam = inp["cam"]
    sc.set_camera(cam.projView, cam.invProjView, cam.m_Pos)
    sc.set_volume(inp["density"])
    sc.set_dir_light((1, 1, 1), -1.57, inp["dir_dir"], 0.0, 1.0)
    sc.set_point_light((0, 0, 0), 1.0, (1, 1, 1))
    sc.set_envmap(inp["env"], None, None, 1.0, 8.0)
    sc.set_mlp(inp["params"], 0.001, 0.0, 0.5)
    sc.set_mrhe(inp["tables"], O.mrhe_resolutions(), 0.01, 0.0, 16, v["table_size"], 2)
    sc.set_train_size(TW, TH)
    sc.set_volume_params(rnd, 1, 0, 0.4, 0.7, 1, 1)
    n_train = TW * TH
    if world == 1:
        idx = np.arange(n_train)
        orec, ocnt = sc.train_trace(TW, TH)
    else:      # a contiguous block of invocations, of which this rank traced every world-th
        b0 = (rank * 977) % max(n_train - 128 * world, 1)
        orec, ocnt = sc.train_trace(TW, TH, b0, b0 + 128 * world)
        sel = np.arange(b0, b0 + 128 * world)
        keep = (sel % world) == rank
        idx, orec, ocnt = sel[keep], orec[keep], ocnt[keep]
    out["train_samples_checked"] = int(idx.size)
    out["train_counters_equal"] = bool(np.array_equal(tcnt[idx][:, :3], ocnt[:, [0, 1, 3]]))
    out["records_bits_equal"] = bool(np.array_equal(rec[idx][:, :6].view(np.uint32), orec[:, :6].view(np.uint32)))
    out["target_rel"] = float(np.abs(rec[idx][:, 6:] - orec[:, 6:]).max() / max(float(np.abs(orec[:, 6:]).max()), 1e-30))
    if world == 1:
        sc.train_accumulate(orec[:, 0:3], orec[:, 3:6], orec[:, 6:9])
        od, ot = sc.get_mlp()[1], sc.get_mrhe()[1][:2 * v["table_size"]]
        out["grad_rel"] = float(np.abs(g_mlp - od).max() / max(float(np.abs(od).max()), 1e-30))
        out["table_grad_rel"] = float(np.abs(g_tab - ot).max() / max(float(np.abs(ot).max()), 1e-30))
        sc.nrc_step(); sc.mrhe_step()
    my_rows = np.arange(rank, H, world)
    step = max(1, int(args.parity_row_step) if world == 1 else -(-len(my_rows) // 48))
    rows = my_rows[step // 2::step]
    eq = {"draws_equal": True, "lit_equal": True, "steps_equal": True, "queries_equal": True, "query_bits_equal": True}
    se, peak, npx = 0.0, 1e-12, 0
    for y in rows:
        y = int(y)
        orgba, oc, oq = sc.render(W, H, y, y + 1)
        eq["draws_equal"] &= bool(np.array_equal(cnt[y, :, 0], oc[y, :, 0]))
        eq["lit_equal"] &= bool(np.array_equal(cnt[y, :, 1], oc[y, :, 1]))
        eq["steps_equal"] &= bool(np.array_equal(cnt[y, :, 2], oc[y, :, 3]))
        eq["queries_equal"] &= bool(np.array_equal(cnt[y, :, 3], oq[y, :, 7].astype(np.uint32)))
        eq["query_bits_equal"] &= bool(np.array_equal(qry[y, :, :6].view(np.uint32), oq[y, :, :6].view(np.uint32)))
        dlt = frame[y].astype(np.float64) - orgba[y].astype(np.float64)
        se += float((dlt * dlt).sum()); npx += dlt.size
        peak = max(peak, float(np.abs(orgba[y]).max()))
    out.update(eq)
    out["rows"] = int(len(rows))
    out["row_step"] = int(step)
    mse = se / max(npx, 1)
    out["frame_psnr_db"] = 99.0 if mse == 0 else float(10.0 * np.log10(peak * peak / mse))
    out["mlp_precision"] = args.precision
    out["note"] = ("GPU frame vs the CPU oracle on the timed scene at its initial weights, random=(0.37,0.61,0.13,0.89); the walk's control flow "
                   "is independent of the MLP precision; frame_psnr_db / grad_rel are for the precision the line times")
    if world > 1:      # all flags must hold on every rank
        flags = [k for k, x in out.items() if isinstance(x, bool)]
        t = torch.tensor([1.0 if out[k] else 0.0 for k in flags] + [-out["frame_psnr_db"]], dtype=torch.float64, device=torch.device("cuda", v["local"]))
        tmin = t.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        for i, k in enumerate(flags):
            out[k] = bool(tmin[i] > 0.5)
        out["frame_psnr_db"] = float(-tmax[-1])      # the worst rank
        out["ranks_checked"] = world
    return out


def finish_b200(args, v):
    import torch
    import torch.distributed as dist
    world, rank, K = v["world"], v["rank"], v["K"]
    tot, tot_e2e, prof = v["tot"], v["tot_e2e"], v["prof"]
    W, H = v["W"], v["H"]
    dev = torch.device("cuda", v["local"])
    # max over ranks of the device time; sums over ranks of the units processed
    t = torch.tensor([v["ms_total"], v["e2e_ms"]], dtype=torch.float64, device=dev)
    c = torch.tensor([tot["infer_queries"] + tot["train_samples"], tot_e2e["infer_queries"] + tot_e2e["train_samples"],
                      v["launches"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c)
    ms_total, e2e_ms = float(t[0]), float(t[1])
    units, units_e2e, launches = float(c[0]), float(c[1]), int(c[2])
    hbm, tflops, which = peaks()
    # dominant kernel of this rank's timed region
    top = max(prof.items(), key=lambda kv: kv[1][0])
    name, (kms, kcount) = top
    my_rows = len(range(rank, H, world)) if world > 1 else H
    bound, work, unit = kernel_work(name, tot, my_rows * W, v["mrhe"].count)
    if name == "k_mrhe_step":
        work = 16.0 * v["mrhe"].count * kcount
    roof = {"kernel": name, "bound": bound, "achieved": None, "peak": hbm if bound == "hbm" else tflops, "unit": unit, "frac": None,
            "traffic": None, "peak_source": which, "launches": kcount, "avg_launch_ms": kms / max(kcount, 1),
            "share_of_step": kms / max(v["ms_total"], 1e-9),
            "note": "share = this kernel's event-timed duration / step time; the two streams overlap, so shares add up to more than 1. "
                    "The walk kernels are latency/issue-bound on an L2-resident grid (profiles/r01p), so their HBM fraction is low by construction: "
                    "'issue_roofline' is the bound that explains the frame time; per-kernel figures for every kernel are under 'kernels'."}
    default_workload = (world == 1 and (W, H, v["TW"], v["TH"]) == (1920, 1080, 100, 100) and args.precision == "bf16")
    if default_workload and name in NCU_FRAME:
        roof["traffic"] = NCU_FRAME[name][0]
        roof["traffic_source"] = "%s (dram__bytes_read.sum + dram__bytes_write.sum, one launch)" % NCU_SOURCE
    if work is not None and kms > 0:
        ach = work / (kms * 1e-3) / (1e9 if bound == "hbm" else 1e12)
        roof["achieved"] = ach
        roof["frac"] = ach / roof["peak"]
        roof["algorithmic_per_launch"] = work / max(kcount, 1)
    kernels = {}
    for kn, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        b, wk, un = kernel_work(kn, tot, my_rows * W, v["mrhe"].count)
        if kn == "k_mrhe_step":
            wk = 16.0 * v["mrhe"].count * cnt
        e = {"ms_per_step": ms / K, "launches_per_step": cnt / K}
        if wk is not None and ms > 0:
            e[un] = wk / (ms * 1e-3) / (1e9 if b == "hbm" else 1e12)
            e["frac_of_peak"] = e[un] / (hbm if b == "hbm" else tflops)
        kernels[kn] = e
    clocks = v["sampler"].summary()
    # What actually bounds the frame: instruction issue.  Warp instructions per frame from the ncu capture over the peak issue rate
    # of the machine at the clock sampled during this run.
    issue = None
    if default_workload and NCU_FRAME:
        winst = sum(x[1] for x in NCU_FRAME.values())
        mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0
        peak_issue = v["ctx"].sm_count * 4 * mhz * 1e6
        issue = {"warp_instructions_per_step": winst, "peak_warp_instructions_per_s": peak_issue,
                 "floor_ms_per_step": winst / peak_issue * 1e3, "frac": (winst / peak_issue * 1e3) / (ms_total / K),
                 "source": "%s (smsp__inst_executed.sum of the frame kernels)" % NCU_SOURCE}
    line = {
        "metric": METRIC, "value": units / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": v["Wm"],
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": v["scaling"], "vs_baseline": None,
        "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
        "config": {"workload": "%s %dx%d 1 spp, online NRC training %dx%d samples/frame, dir+point+env lights, useNN=1" % (
                       "cloud_sixteenth (synthetic 125x85x153 stand-in)" if v["grid_scale"] == 1 else
                       "full cloud (synthetic 125x85x153 grid refined x%d per axis)" % v["grid_scale"], W, H, v["TW"], v["TH"]),
                   "baseline_config": args.config,
                   "envmap": ("file " + os.path.basename(os.environ["NRC_ENVMAP"])) if os.environ.get("NRC_ENVMAP") else "synthetic sky 1024x512 (the shape of data/image/photostudio.hdr; set NRC_ENVMAP=<file.hdr> to load it)",
                   "frame": [W, H], "train_grid": [v["TW"], v["TH"]], "hash_table": [16, v["table_size"], 2], "mlp": "64x5->3 relu",
                   "parallelism": "rows interleaved over %d GPU(s), train samples sharded, %s" % (
                       world, "gradients summed over NVLink peer memory inside the fused step kernel (sharded form for heads above 2^20 floats)" if (world > 1 and args.exchange == "fused")
                       else "NCCL all-reduce of gradients"),
                   "vertex_pool": vertex_pool_note(v),
                   "l2": "fresh per-frame RNG seed: every frame traces new paths; per-frame buffers (%.0f MB) exceed nothing but are rewritten, "
                         "the density grid (%s) %s" % (W * H * 120 / 1e6, "13 MB of bricks" if v["grid_scale"] == 1 else "838 MB of bricks",
                                                      "is L2-resident by design" if v["grid_scale"] == 1 else "does not fit L2: taps gather from HBM")},
        "e2e": {"value": units_e2e / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / K,
                "h2d_bytes_per_step": 40 + 140 + 36 + 28 + 8, "d2h_bytes_per_step": v["d2h_bytes"] + 4},
        "gpu_launches": launches, "roofline": roof, "issue_roofline": issue, "kernels": kernels, "clocks": clocks,
        "queries_per_step": {"infer": tot["infer_queries"] / K, "train": tot["train_samples"] / K,
                             "density_taps": (tot["render_taps"] + tot["train_taps"]) / K},
        "mse_loss_last": v["mse"],
    }
    if v.get("fp32_ms") is not None:
        line["fp32_mlp_path"] = {"ms_per_step": v["fp32_ms"], "note": "same frames with the fp32 CUDA-core MLP kernels (1e-3 parity path), this rank"}
    if world == 1 and not args.no_fit:
        try:
            import neuralradiancecaching_b200 as N
            from neuralradiancecaching_b200 import _lib as L
            line["fit_config2"] = run_fit(v["ctx"], N, L, torch)
        except Exception as e:
            line["fit_config2"] = {"error": repr(e)}
        try:
            line["mlp_stream"] = run_mlp_stream(v["ctx"], N, torch)
        except Exception as e:
            line["mlp_stream"] = {"error": repr(e)}
    if world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_reference_sample(W, H, v["TW"], v["TH"], budget_s=args.cpu_budget)
        except Exception as e:   # the checker failing must not hide the GPU number
            line["cpu_baseline"] = {"error": repr(e)}
    if not args.no_side and not args.no_parity:
        try:
            line["parity"] = parity_check(v, args)
        except Exception as e:   # the checker failing must not hide the GPU number
            line["parity"] = {"error": repr(e)}
    line["repeats"] = {"count": len(v["rep_ms"]), "steps_each": K, "ms_per_step": [x / K for x in v["rep_ms"]],
                       "e2e_ms_per_step": [x / K for x in v["e2e_rep_ms"]], "reported": "median repeat (max over ranks per repeat)"}
    v["wl"].destroy()
    # ---- the other BASELINE configurations, measured briefly behind the headline so that they appear in driver-run lines ----
    if not args.no_side and args.config == "default" and not args.width:
        ctx_, local = v["ctx"], v["local"]
        side = []
        if world == 1:
            side.append(("config3", (1280, 720, 100, 100, 1, 16384, "none"), "BASELINE configs[2]: 1280x720, 1 spp, online training, env map, 1 GPU"))
            side.append(("strong_4k", (3840, 2160, 100, 100, 1, 16384, "strong"),
                         "BASELINE configs[3] on one GPU: the 3840x2160 frame + 10^4 samples whose split over 2/4/8 ranks is the strong-scaling series"))
            side.append(("config5_n1", (1920, 1080, 100, 100, 4, 1 << 22, "none"),
                         "BASELINE configs[4] on one GPU: refined 500x340x612 cloud, 2^22-entry hash tables, 1080p"))
        else:
            side.append(("strong_4k", (3840, 2160, 100, 100, 1, 16384, "strong"),
                         "BASELINE configs[3]: ONE 3840x2160 frame + 10^4 samples split over the ranks (strong scaling)"))
            if world >= 8:
                Ws, Hs, TWs, THs = frame_shape(world)
                side.append(("config5", (Ws, Hs, TWs, THs, 4, 1 << 22, "weak"),
                             "BASELINE configs[4]: refined cloud, 2^22-entry hash tables, %d GPUs (weak-scaling frame)" % world))
        for key, (w_, h_, tw_, th_, gs_, ts_, sc_), label in side:
            if key in (os.environ.get("NRC_BENCH_SKIP", "") or "").split(","):
                continue
            try:
                line[key] = side_measure(args, ctx_, rank, world, local, w_, h_, tw_, th_, gs_, ts_, sc_, label=label)
            except SystemExit:
                raise
            except Exception as e:
                line[key] = {"error": repr(e)}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


# ---- the CPU arm: the oracle (a C++ restatement of the reference's shaders; the reference has no CPU code for this path) ------
def cpu_reference_sample(W, H, TW, TH, budget_s=20.0, steps=1, rows_per_step=None):
    """Times the oracle on a bounded, evenly spread sample of the same frame: every (H/rows)-th frame row through the
    render pass and evenly spread chunks of the train grid through trace + Backprop, then both optimiser steps.
    value = (inference queries + training samples processed) / wall time, on all host threads."""
    from oracle import oracle as O
    inp = scene_inputs(W, H)
    sc = O.OracleScene()
    cam = inp["cam"]
    sc.set_camera(cam.projView, cam.invProjView, cam.m_Pos)
    sc.set_volume(inp["density"])
    sc.set_dir_light((1, 1, 1), -1.57, inp["dir_dir"], 0.0, 1.0)
    sc.set_point_light((0, 0, 0), 1.0, (1, 1, 1))
    sc.set_envmap(inp["env"], None, None, 1.0, 8.0)
    sc.set_mlp(inp["params"], 0.001, 0.0, 0.5)
    sc.set_mrhe(inp["tables"], O.mrhe_resolutions(), 0.01, 0.0, 16, 16384, 2)
    sc.set_train_size(TW, TH)
    rnd = lcg_randoms(steps + 1)
    # all host threads: torchrun exports OMP_NUM_THREADS=1 to its workers, which would time a one-thread baseline
    want_threads = int(os.environ.get("NRC_CPU_THREADS", "0")) or len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    O.set_num_threads(want_threads)
    threads = O.num_threads()
    import ctypes as C
    rgba = np.zeros((H, W, 4), np.float32)
    qry = np.zeros((H, W, 8), np.float32)

    def one(rows, chunks, chunk, random4):
        sc.set_volume_params(random4, 1, 0, 0.4, 0.7, 1, 1)
        q = 0
        t0 = time.perf_counter()
        n_train = TW * TH
        recs = []
        for k in range(chunks):                       # train: trace evenly spread chunks of the grid ...
            i0 = (k * n_train) // chunks
            recs.append(sc.train_trace(TW, TH, i0, min(i0 + chunk, n_train))[0])
        rec = np.concatenate(recs)                    # ... and Backprop them as one batch (all threads busy)
        sc.train_accumulate(rec[:, 0:3], rec[:, 3:6], rec[:, 6:9])
        q += rec.shape[0]
        sc.nrc_step()
        sc.mrhe_step()
        for k in range(rows):                         # render: evenly spread rows
            y = (k * H) // rows + (H // rows) // 2
            sc.l.orc_render(sc.h, C.c_uint(W), C.c_uint(H), C.c_uint(y), C.c_uint(y + 1), O._p(rgba), None, O._p(qry))
            q += int(qry[y, :, 7].sum())
        return q, time.perf_counter() - t0

    # calibrate on a tiny sample, then size the sample for the budget; rows and train samples keep the frame's proportions
    n_train = TW * TH
    chunks = 25
    q0, t0 = one(4, 4, max(1, -(-n_train * 4 // H) // 4), rnd[0])
    per_row = max(t0 / 4.0, 1e-4)
    rows = rows_per_step or int(max(4, min(H, budget_s / per_row / max(steps, 1))))
    chunk = int(max(1, min(n_train // chunks, -(-n_train * rows // H) // chunks)))
    total_q, total_t = 0, 0.0
    for s in range(steps):
        q, t = one(rows, chunks, chunk, rnd[s + 1])
        total_q += q
        total_t += t
    return {"value": total_q / total_t, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d of %d frame rows (evenly spread) + %d of %d train samples (%d chunks) per step, %d step(s), %.1f s; "
                      "oracle = C++ restatement of the reference's GLSL (no CPU implementation exists in the reference; "
                      "lavapipe unavailable on this box)" % (rows, H, chunks * chunk, TW * TH, chunks, steps, total_t),
            "seconds": total_t, "queries": total_q}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    W, H, TW, TH = frame_shape(world)
    K, Wm = args.steps, args.warmup
    # each "step" is a bounded sample of one frame; the whole run stays within a few minutes
    per_step_budget = max(2.0, min(20.0, 120.0 / max(K + Wm, 1)))
    cpu_reference_sample(W, H, TW, TH, budget_s=2.0, steps=max(min(Wm, 1), 0) or 1)      # warm caches, page in
    t0 = time.perf_counter()
    cb = cpu_reference_sample(W, H, TW, TH, budget_s=per_step_budget * K, steps=K)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": cb["seconds"] / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "cloud_sixteenth (synthetic 125x85x153 stand-in) %dx%d 1 spp, online NRC training %dx%d samples/frame, "
                                   "dir+point+env lights, useNN=1" % (W, H, TW, TH),
                       "baseline_config": "default", "frame": [W, H], "train_grid": [TW, TH], "hash_table": [16, 16384, 2], "mlp": "64x5->3 relu",
                       "parallelism": "CPU oracle on the host cores of rank 0: a bounded, evenly spread sample of each frame (see cpu_baseline.sample)"},
            "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"],
                    help="MLP arithmetic: bf16 = tcgen05 tensor cores (bf16 operands, fp32 accumulate; north_star's tensor-core path), "
                         "fp32 = CUDA-core 1e-3 parity path.  Tracing is fp32 either way.")
    ap.add_argument("--no-fp32-line", action="store_true", help="skip the extra fp32-path measurement")
    ap.add_argument("--exchange", default=os.environ.get("NRC_EXCHANGE", "fused"), choices=["nccl", "fused"],
                    help="multi-GPU gradient exchange: two NCCL all-reduces + step kernels, or the fused peer-memory kernel (csrc/dp.cu)")
    ap.add_argument("--config", default="default", choices=["default", "3", "4", "5"],
                    help="BASELINE.json configs[i] other than the headline one: 3 = 1280x720 on one GPU, 4 = one 3840x2160 frame split "
                         "over the ranks (strong scaling), 5 = refined cloud grid + 2^22-entry hash tables")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fit", action="store_true", help="skip the BASELINE config-2 (2^18-query MRHE+MLP fit) side measurement")
    ap.add_argument("--no-side", action="store_true", help="skip the side measurements of the other BASELINE configs (config3 / strong_4k / config5) and the parity check")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--repeats", type=int, default=3, help="the K timed steps are repeated this many times; the median repeat is reported")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle parity check of the timed workload")
    ap.add_argument("--parity-row-step", type=int, default=4, help="N = 1: the parity check compares every k-th frame row with the oracle")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
