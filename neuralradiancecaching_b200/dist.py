"""Data-parallel plumbing: one process per GPU, torch.distributed for the rendezvous and the collective.

The reference has no multi-GPU path; the protocol is ours:
  rows / train invocations are interleaved over ranks (rank, rank+world, ...) with GLOBAL coordinates in the RNG seeds,
  gradients (packed MLP arena incl. the loss accumulator + the touched head of the delta hash table) are summed over
  ranks between nrc-train and the optimiser steps, every rank then applies identical steps."""
import numpy as np


class DevArray:
    """__cuda_array_interface__ view of a raw device pointer (torch.as_tensor makes a zero-copy tensor of it)."""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def strided_count(first, end, stride):
    """number of indices first, first+stride, ... < end"""
    return 0 if first >= end else (end - first + stride - 1) // stride


def shard_rows(rank, world, height):
    return np.arange(rank, height, world)


def frame_shape(n_gpus, base=(1920, 1080, 100, 100)):
    """Weak scaling: n x the pixels of the base frame (aspect kept, both sides a multiple of 8) and n x the samples."""
    w, h, tw, th = base
    s = float(n_gpus) ** 0.5
    return int(round(w * s / 8.0)) * 8, int(round(h * s / 8.0)) * 8, tw, th * n_gpus


class GradientExchange:
    """All-reduce of the two gradient arenas, in place, ordered on the library's stream."""

    def __init__(self, ctx, nrc, mrhe, device):
        import torch
        gm_ptr, gm_n = nrc.grad_ptr()
        gt_ptr, gt_n = mrhe.grad_ptr()
        self.g_mlp = torch.as_tensor(DevArray(gm_ptr, gm_n), device=device)
        self.g_tab = torch.as_tensor(DevArray(gt_ptr, gt_n), device=device)
        self.stream = torch.cuda.ExternalStream(ctx.stream, device=device)

    def __call__(self):
        import torch
        import torch.distributed as dist
        with torch.cuda.stream(self.stream):
            dist.all_reduce(self.g_mlp)
            dist.all_reduce(self.g_tab)


class FusedExchange:
    """Peer-memory exchange fused with both optimiser steps (csrc/dp.cu): one kernel per rank and frame instead of two
    NCCL all-reduces + two step kernels.  __call__ replaces `exchange(); nrc.Step(); mrhe.Step()`.

    handles=None: multi-process; the IPC handles are gathered with torch.distributed (any backend).
    handles=[...] : replicas inside this process (tests), see connect_local()."""

    def __init__(self, nrc, mrhe, rank, world, device=None, connect=True):
        import ctypes as C
        from . import _lib as L
        self._L, self._C = L, C
        self.rank, self.world = rank, world
        self.h = C.c_void_p()
        L.check(L.lib().nrc_dp_create(nrc.h, mrhe.h, C.c_int(rank), C.c_int(world), C.byref(self.h)))
        n = int(L.lib().nrc_dp_handle_bytes())
        buf = (C.c_ubyte * n)()
        L.check(L.lib().nrc_dp_export(self.h, buf))
        self.handle = bytes(buf)
        if connect:
            self._connect_distributed(device)

    def _connect_distributed(self, device):
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return self.connect([self.handle], same_process=True)
        on_gpu = dist.get_backend() == "nccl"
        mine = torch.tensor(list(self.handle), dtype=torch.uint8, device=device if on_gpu else "cpu")
        parts = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(parts, mine)
        self.connect([bytes(p.cpu().numpy().tobytes()) for p in parts], same_process=False)
        dist.barrier()   # nobody launches before every rank has mapped its peers

    def connect(self, handles, same_process):
        C, L = self._C, self._L
        blob = b"".join(handles)
        arr = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        L.check(L.lib().nrc_dp_connect(self.h, arr, C.c_int(1 if same_process else 0)))

    def __call__(self):
        self._L.check(self._L.lib().nrc_dp_reduce_step(self.h))

    def phase(self, k):
        self._L.check(self._L.lib().nrc_dp_reduce_step_phase(self.h, self._C.c_int(k)))

    def timed_out(self):
        """True if a handshake ever gave up waiting for a peer (waits for the stream)."""
        v = self._C.c_int(0)
        self._L.check(self._L.lib().nrc_dp_status(self.h, self._C.byref(v)))
        return bool(v.value)

    def close(self):
        if self.h:
            self._L.check(self._L.lib().nrc_dp_destroy(self.h))
            self.h = None


def gather_rows(part, height, rank, world, dst=0):
    """Frame assembly: every rank holds its interleaved rows (rank, rank+world, ...) compacted in `part`
    ([rows, width, channels] torch tensor, CPU for gloo or CUDA for NCCL); rank `dst` gets the [height, width, channels] frame,
    the others None.  Row counts differ by at most one between ranks: parts are padded to the longest."""
    import torch
    import torch.distributed as dist
    rows_max = strided_count(0, height, world)
    pad = part
    if part.shape[0] < rows_max:
        pad = torch.zeros((rows_max,) + tuple(part.shape[1:]), dtype=part.dtype, device=part.device)
        pad[:part.shape[0]] = part
    pad = pad.contiguous()
    if world == 1:
        return pad[:height]
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)            # 4 B/pixel at BGRA8: small next to a frame's work; gather-to-one is not in every backend
    if rank != dst:
        return None
    full = torch.empty((height,) + tuple(part.shape[1:]), dtype=part.dtype, device=part.device)
    for k in range(world):
        n = strided_count(k, height, world)
        full[k::world] = parts[k][:n]
    return full


def allreduce_host(arrays):
    """gloo/CPU variant used by the world_size-2 tests: numpy arrays summed over ranks in place."""
    import torch
    import torch.distributed as dist
    for a in arrays:
        t = torch.from_numpy(a)
        dist.all_reduce(t)
    return arrays
