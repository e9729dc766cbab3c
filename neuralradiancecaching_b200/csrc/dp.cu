// dp.cu -- data-parallel gradient exchange fused with the optimiser steps, over NVLink peer memory.
//
// What the NCCL path does with four launches per frame (all-reduce of the MLP gradient arena, all-reduce of the touched
// head of the delta hash table, nrc-step, mrhe-step) this does with ONE kernel per rank: every rank maps its peers'
// gradient arenas (CUDA IPC), and the kernel
//   1. tells every peer "my gradients are complete" and waits for the same from all of them (flags in peer memory,
//      release/acquire at system scope),
//   2. for each parameter reads the W gradients straight from the W GPUs (ld.relaxed.sys over NVLink), adds them in RANK
//      ORDER -- so every replica computes bit-identical sums, which a ring all-reduce does not promise -- and applies
//      nrc-step.comp / mrhe-step.comp to its own replica in the same thread,
//   3. tells every peer "I have read your gradients", waits for all, and clears its own arenas.
// The exchange is 84 KB + 2*T*4 B: latency-bound (the two NCCL all-reduces alone measure 60 us at 4 GPUs), so the kernel is
// 16 small blocks that each own a fixed slice of both arenas and handshake with the same block of the peers -- small
// blocks get an SM slot at once while the render chain fills the GPU, and nothing needs a grid-wide barrier.
// Bandwidth regime (heads above NRC_DP_SHARD_FLOATS: T = 2^22 tables, 33.5 MB; the fixed preset, 512 MiB): the same kernel in
// SHARDED form -- reduce-scatter + sharded step + all-gather in one launch.  Rank r owns slice r of the head: it reads the W gradient
// slices over NVLink, adds them in rank order, applies mrhe-step.comp (or the opt-in Adam) to its slice and WRITES the new table
// values (and Adam moments) into every replica (st.relaxed.sys.v4).  Per rank 2 x (W-1)/W x head bytes cross NVLink instead of
// (W-1) x head; the grid grows to one block per SM.  The handshakes are the same two: "gradients complete" before, "I have read yours and
// written my slice of your table" after -- the second one is what makes the peers' table writes visible before the next kernel here.
// The reference has no multi-GPU path at all; the step arithmetic is the one of steps.cu (nrc-step.comp:200-406,
// mrhe-step.comp:27-74).
#include <cuda_bf16.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>
#include "nrc_internal.cuh"
#include "tc_ptx.cuh"

#define NRC_DP_MAX_WORLD 16
#define NRC_DP_SHARD_FLOATS (1u << 20)   // heads above this many floats are exchanged in sharded form (NRC_DP_SHARD=1 forces it)
#define NRC_DP_MAX_BLOCKS 592   // upper bound of the grid (flag slots); the sharded form uses one block per SM by default
#define NRC_DP_FLAG_COUNT (NRC_DP_MAX_BLOCKS * NRC_DP_MAX_WORLD * 2)

struct nrc_dp {
    nrc_mlp* mlp = nullptr;
    nrc_mrhe* mrhe = nullptr;
    int rank = 0, world = 1;
    unsigned* flags = nullptr;                 // [DP_BLOCKS][NRC_DP_MAX_WORLD][2], written by the peers
    unsigned epoch = 0;
    bool connected = false, sameProcess = false;
    const float* peerGrad[NRC_DP_MAX_WORLD] = {nullptr};
    const float* peerDelta[NRC_DP_MAX_WORLD] = {nullptr};
    unsigned* peerFlags[NRC_DP_MAX_WORLD] = {nullptr};
    float* peerTables[NRC_DP_MAX_WORLD] = {nullptr};   // sharded form: the owner of a slice writes it into every replica
    float* peerM1[NRC_DP_MAX_WORLD] = {nullptr};
    float* peerM2[NRC_DP_MAX_WORLD] = {nullptr};
    void* opened[NRC_DP_MAX_WORLD][6] = {{nullptr}};
    bool sharded = false;
    bool mlpAdam = false, mrheAdam = false;    // as they were when the exchange was created (the moments' buffers are mapped then)
    unsigned* hostFail = nullptr;              // pinned mirror of the sticky time-out word, refreshed behind every launch
};

struct DpHandle {   // what one rank publishes to the others
    cudaIpcMemHandle_t grad, delta, flags, tables, m1, m2;
    unsigned long long rawGrad, rawDelta, rawFlags, rawTables, rawM1, rawM2;   // same-process emulation
    unsigned long long extent;
    int rank, adam;
};

struct DpArgs {
    const float* peerGrad[NRC_DP_MAX_WORLD];
    const float* peerDelta[NRC_DP_MAX_WORLD];
    unsigned* peerFlags[NRC_DP_MAX_WORLD];
    unsigned* myFlags;
    int rank, world;
    unsigned epoch;
    int barriers, clear, doStep;   // clear: 1 = clear arenas + publish the global loss, 2 = clear arenas only
    float* params; float* grad; float* momentum; float* paramsT; unsigned short* imgBf16;
    float lr, wd, beta1;
    float* tables; float* delta; unsigned extent;
    float mlr, mwd;
    // sharded form: destinations of the owner's writes; nDest = world (sharded) or 1 (replicated: this replica only)
    float* destTables[NRC_DP_MAX_WORLD]; float* destM1[NRC_DP_MAX_WORLD]; float* destM2[NRC_DP_MAX_WORLD];
    int nDest, sharded;
    // opt-in Adam (steps.cu: k_nrc_step_adam / k_mrhe_step_adam): moments of this replica, bias corrections computed on the host
    int mlpAdam, mrheAdam;
    float* second; float beta2, eps, c1, c2;
    float* m1; float* m2; float mbeta1, mbeta2, meps, mc1, mc2;
    unsigned long long timeoutNs;   // how long a handshake waits for a peer before the exchange fails closed
};

__device__ __forceinline__ float4 ld_sys4(const float* p)
{
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p)
{
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ bool dp_bad(float x) { return (x != x) || (fabsf(x) == INFINITY) || (fabsf(x) > 1000.0f); }
__device__ __forceinline__ float dp_sign(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }

// Block b of every rank owns the same slice of both arenas, so the handshakes are per block: block b signals slot
// (b, rank, which) of every peer and waits until block b of every peer has signalled its own.  No grid-wide barrier.
// A peer that never arrives (crashed process) must not hang this GPU: the wait gives up after ~30 s and the exchange FAILS CLOSED --
// s_fail is raised, the caller skips its step and its clear for this launch, and the sticky word NRC_DP_FLAG_COUNT is set in this
// rank's flag block AND in every peer's, so that every rank's next nrc_dp_reduce_step / nrc_sync / nrc_readback_wait reports it.
// A block that has already failed only signals (its peers must not wait out their own time-out for it) and does not wait.
__device__ __forceinline__ void dp_barrier(const DpArgs& a, int which, int* s_fail)
{
    __syncthreads();
    if ((int)threadIdx.x < a.world) {
        __threadfence_system();
        unsigned slot = (blockIdx.x * NRC_DP_MAX_WORLD) * 2;
        st_release_sys(a.peerFlags[threadIdx.x] + slot + a.rank * 2 + which, a.epoch);
        if (!*s_fail) {
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while ((int)(ld_acquire_sys(a.myFlags + slot + threadIdx.x * 2 + which) - a.epoch) < 0) {
                __nanosleep(32);
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > a.timeoutNs) {
                    atomicExch(s_fail, 1);
                    for (int r = 0; r < a.world; r++) st_release_sys(a.peerFlags[r] + NRC_DP_FLAG_COUNT, 1u);
                    break;
                }
            }
        }
    }
    __syncthreads();
}

// sum of one float4 over the ranks, in rank order; all W loads are in flight before the first add
__device__ __forceinline__ float4 dp_sum4(const float* const* peers, int world, size_t i4)
{
    float4 v[NRC_DP_MAX_WORLD];
#pragma unroll
    for (int r = 0; r < NRC_DP_MAX_WORLD; r++)
        if (r < world) v[r] = ld_sys4(peers[r] + i4 * 4);
    float4 g = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
    for (int r = 0; r < NRC_DP_MAX_WORLD; r++)
        if (r < world) { g.x = __fadd_rn(g.x, v[r].x); g.y = __fadd_rn(g.y, v[r].y); g.z = __fadd_rn(g.z, v[r].z); g.w = __fadd_rn(g.w, v[r].w); }
    return g;
}

__device__ __forceinline__ void dp_nrc_update(const DpArgs& a, int i, float g)
{
    float w = a.params[i];
    if (dp_bad(g)) g = __fmul_rn(1000.0f, dp_sign(g));
    g = __fadd_rn(g, __fmul_rn(-w, a.wd));
    float d = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, a.beta1), g), __fmul_rn(a.beta1, a.momentum[i]));
    a.momentum[i] = d;
    if (a.mlpAdam) {   // k_nrc_step_adam, operation for operation
        float v = __fadd_rn(__fmul_rn(a.beta2, a.second[i]), __fmul_rn(__fsub_rn(1.0f, a.beta2), __fmul_rn(g, g)));
        a.second[i] = v;
        d = __fdiv_rn(__fmul_rn(d, a.c1), __fadd_rn(__fsqrt_rn(__fmul_rn(v, a.c2)), a.eps));
    }
    float nw = __fadd_rn(w, __fmul_rn(d, a.lr));
    if (i < NRC_MLP_WEIGHT_COUNT && dp_bad(nw)) nw = 0.0f;
    a.params[i] = nw;
    if (i < 5 * 4096) {
        int l = i >> 12, rem = i & 4095, o = rem >> 6, k = rem & 63;
        a.paramsT[l * 4096 + k * 64 + o] = nw;
    }
    if (a.imgBf16 && i < NRC_MLP_WEIGHT_COUNT) tc_store_weight(a.imgBf16, i, nw);   // the operand images of the tcgen05 path, as k_nrc_step
}

__device__ __forceinline__ float dp_mrhe_update(float w, float g, float lr, float wd)
{
    if (dp_bad(g)) g = __fmul_rn(100.0f, dp_sign(g));
    g = __fadd_rn(g, __fmul_rn(-w, wd));
    float nw = __fadd_rn(w, __fmul_rn(g, lr));
    return dp_bad(nw) ? 0.0f : nw;
}

// mrhe_update_adam of steps.cu, operation for operation
__device__ __forceinline__ float dp_mrhe_update_adam(const DpArgs& a, float w, float g, float& m1, float& m2)
{
    if (dp_bad(g)) g = __fmul_rn(100.0f, dp_sign(g));
    g = __fadd_rn(g, __fmul_rn(-w, a.mwd));
    m1 = __fadd_rn(__fmul_rn(a.mbeta1, m1), __fmul_rn(__fsub_rn(1.0f, a.mbeta1), g));
    m2 = __fadd_rn(__fmul_rn(a.mbeta2, m2), __fmul_rn(__fsub_rn(1.0f, a.mbeta2), __fmul_rn(g, g)));
    float d = __fdiv_rn(__fmul_rn(m1, a.mc1), __fadd_rn(__fsqrt_rn(__fmul_rn(m2, a.mc2)), a.meps));
    float nw = __fadd_rn(w, __fmul_rn(d, a.mlr));
    return dp_bad(nw) ? 0.0f : nw;
}

__device__ __forceinline__ void st_sys4(float* p, float4 v)
{
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

#define DP_BLOCKS 16
#define DP_THREADS 256
#define DP_MLP_VEC ((NRC_MLP_PARAM_COUNT + 1 + 3) / 4)   // 5249 float4: 20995 gradients + the loss accumulator

__global__ void __launch_bounds__(DP_THREADS) k_dp_reduce_step(const __grid_constant__ DpArgs a)
{
    const int t = blockIdx.x * DP_THREADS + threadIdx.x;
    const int stride = gridDim.x * DP_THREADS;
    __shared__ int s_fail;
    if (threadIdx.x == 0) s_fail = a.barriers ? (ld_acquire_sys(a.myFlags + NRC_DP_FLAG_COUNT) != 0u) : 0;   // an earlier launch (here or on a peer) gave up
    __syncthreads();
    if (a.barriers) dp_barrier(a, 0, &s_fail);
    float lossSum = 0.0f;
    if (a.doStep && !s_fail) {
        // ---- MLP arena: sum over ranks in rank order, then nrc-step.comp for this replica ----
        for (int i4 = t; i4 < DP_MLP_VEC; i4 += stride) {
            float4 g = dp_sum4(a.peerGrad, a.world, i4);
            float gs[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
            for (int c = 0; c < 4; c++) {
                int i = i4 * 4 + c;
                if (i < NRC_MLP_PARAM_COUNT) dp_nrc_update(a, i, gs[c]);
                else if (i == NRC_MLP_PARAM_COUNT) lossSum = gs[c];
            }
        }
        // ---- head of the delta hash table (all Backprop can touch): sum + mrhe-step.comp.  Replicated form: every rank
        // reduces and steps the whole head for itself.  Sharded form: this rank's slice only, written into every replica. ----
        const unsigned n4 = a.extent / 4;
        unsigned lo = 0, hi = n4;
        if (a.sharded) {
            const unsigned per = (n4 + (unsigned)a.world - 1u) / (unsigned)a.world;
            lo = min(per * (unsigned)a.rank, n4);
            hi = min(lo + per, n4);
        }
#pragma unroll 2
        for (unsigned i4 = lo + t; i4 < hi; i4 += stride) {
            float4 g = dp_sum4(a.peerDelta, a.world, i4);
            float4 w = reinterpret_cast<const float4*>(a.tables)[i4];
            if (a.mrheAdam) {
                float4 p = reinterpret_cast<const float4*>(a.m1)[i4], q = reinterpret_cast<const float4*>(a.m2)[i4];
                w.x = dp_mrhe_update_adam(a, w.x, g.x, p.x, q.x);
                w.y = dp_mrhe_update_adam(a, w.y, g.y, p.y, q.y);
                w.z = dp_mrhe_update_adam(a, w.z, g.z, p.z, q.z);
                w.w = dp_mrhe_update_adam(a, w.w, g.w, p.w, q.w);
                for (int r = 0; r < a.nDest; r++) { st_sys4(a.destM1[r] + (size_t)i4 * 4, p); st_sys4(a.destM2[r] + (size_t)i4 * 4, q); }
            } else {
                w.x = dp_mrhe_update(w.x, g.x, a.mlr, a.mwd);
                w.y = dp_mrhe_update(w.y, g.y, a.mlr, a.mwd);
                w.z = dp_mrhe_update(w.z, g.z, a.mlr, a.mwd);
                w.w = dp_mrhe_update(w.w, g.w, a.mlr, a.mwd);
            }
            for (int r = 0; r < a.nDest; r++) st_sys4(a.destTables[r] + (size_t)i4 * 4, w);
        }
    }
    if (a.barriers) dp_barrier(a, 1, &s_fail);     // block b of every peer has read my slice b (and, sharded, written its slice of my table)
    if (a.clear && !s_fail) {
        for (int i4 = t; i4 < DP_MLP_VEC; i4 += stride) {
            float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (i4 == DP_MLP_VEC - 1) {   // {b5[0..2] gradients, loss accumulator}
                // The exchanged slot [20995] holds what THIS rank accumulated since the last exchange; it restarts at zero.  The global
                // sum goes to the running total in slot [20996], which is not exchanged -- GetStats() reports [20995] + [20996], and
                // ResetStats() clears both.  (Round 1 wrote the global sum back into the exchanged slot: a host that did not reset the
                // statistics every frame saw W * previous + new after the next exchange.)
                float own = a.grad[NRC_MLP_PARAM_COUNT];
                if (a.clear == 1) a.grad[NRC_MLP_PARAM_COUNT + 1] += lossSum;
                z.w = (a.clear == 1) ? 0.0f : own;
            }
            reinterpret_cast<float4*>(a.grad)[i4] = z;
        }
        // Block b may only clear what block b of the READER has read (the handshakes pair equal block indices): in the sharded form slice s
        // of my gradients is read by rank s, whose block b walks lo_s + t, lo_s + t + stride, ... -- so the clear walks every slice the same way.
        const unsigned n4 = a.extent / 4;
        const unsigned per = a.sharded ? (n4 + (unsigned)a.world - 1u) / (unsigned)a.world : n4;
        for (unsigned lo = 0; lo < n4; lo += per) {
            const unsigned hi = min(lo + per, n4);
            for (unsigned i4 = lo + t; i4 < hi; i4 += stride) reinterpret_cast<float4*>(a.delta)[i4] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
    }
}

// ---- host side -----------------------------------------------------------------------------------------------------------
nrc_status launch_mrhe_step_range(nrc_mrhe* m, size_t firstFloat, cudaStream_t st);   // steps.cu

extern "C" size_t nrc_dp_handle_bytes(void) { return sizeof(DpHandle); }

extern "C" nrc_status nrc_dp_create(nrc_mlp* mlp, nrc_mrhe* mrhe, int rank, int world, nrc_dp** out)
{
    NRC_REQUIRE(mlp && mrhe && out, "null argument");
    NRC_REQUIRE(world >= 1 && world <= NRC_DP_MAX_WORLD && rank >= 0 && rank < world, "rank/world out of range");
    NRC_REQUIRE(nrc_mrhe_grad_extent(mrhe) < ((size_t)1 << 32) && (nrc_mrhe_grad_extent(mrhe) % 4) == 0, "delta-table head out of range for the fused exchange");
    nrc_dp* d = new nrc_dp();
    d->mlp = mlp; d->mrhe = mrhe; d->rank = rank; d->world = world;
    {
        const char* e = getenv("NRC_DP_SHARD");   // 1 / 0 force the sharded / replicated form (tests, measurements)
        d->sharded = e ? (atoi(e) != 0) : (nrc_mrhe_grad_extent(mrhe) > NRC_DP_SHARD_FLOATS);
        if (world == 1) d->sharded = false;
    }
    d->mlpAdam = mlp->adam; d->mrheAdam = mrhe->adam;
    mlp->lossTotalSlot = true;   // GetStats() must add the running total the exchange keeps in grad[20996]
    cudaError_t e = cudaMalloc(&d->flags, sizeof(unsigned) * (NRC_DP_FLAG_COUNT + 1));
    if (e == cudaSuccess) e = cudaMemset(d->flags, 0, sizeof(unsigned) * (NRC_DP_FLAG_COUNT + 1));
    if (e != cudaSuccess) {
        nrc_set_error("nrc_dp_create: %s", cudaGetErrorString(e));
        delete d;
        return NRC_ERR_CUDA;
    }
    if (cudaMallocHost(&d->hostFail, sizeof(unsigned)) != cudaSuccess) {
        nrc_set_error("nrc_dp_create: cudaMallocHost failed");
        cudaFree(d->flags);
        delete d;
        return NRC_ERR_CUDA;
    }
    *d->hostFail = 0u;
    mlp->ctx->dpFailWords.push_back(d->hostFail);
    *out = d;
    return NRC_OK;
}

extern "C" nrc_status nrc_dp_export(nrc_dp* d, void* handle)
{
    NRC_REQUIRE(d && handle, "null argument");
    DpHandle h;
    memset(&h, 0, sizeof(h));
    NRC_CUDA(cudaIpcGetMemHandle(&h.grad, d->mlp->grad));
    NRC_CUDA(cudaIpcGetMemHandle(&h.delta, d->mrhe->delta));
    NRC_CUDA(cudaIpcGetMemHandle(&h.flags, d->flags));
    h.rawGrad = (unsigned long long)(uintptr_t)d->mlp->grad;
    h.rawDelta = (unsigned long long)(uintptr_t)d->mrhe->delta;
    h.rawFlags = (unsigned long long)(uintptr_t)d->flags;
    NRC_CUDA(cudaIpcGetMemHandle(&h.tables, d->mrhe->tables));
    h.rawTables = (unsigned long long)(uintptr_t)d->mrhe->tables;
    if (d->mrheAdam) {
        NRC_CUDA(cudaIpcGetMemHandle(&h.m1, d->mrhe->m1));
        NRC_CUDA(cudaIpcGetMemHandle(&h.m2, d->mrhe->m2));
        h.rawM1 = (unsigned long long)(uintptr_t)d->mrhe->m1;
        h.rawM2 = (unsigned long long)(uintptr_t)d->mrhe->m2;
    }
    h.extent = nrc_mrhe_grad_extent(d->mrhe);
    h.rank = d->rank;
    h.adam = (d->mlpAdam ? 1 : 0) | (d->mrheAdam ? 2 : 0) | (d->sharded ? 4 : 0);
    memcpy(handle, &h, sizeof(h));
    return NRC_OK;
}

// allHandles: world handles in rank order.  sameProcess != 0: the "ranks" are replicas inside this process (raw device
// pointers are used; the single-GPU parity test runs the exchange that way).
extern "C" nrc_status nrc_dp_connect(nrc_dp* d, const void* allHandles, int sameProcess)
{
    NRC_REQUIRE(d && allHandles, "null argument");
    NRC_REQUIRE(!d->connected, "already connected");
    const DpHandle* hs = static_cast<const DpHandle*>(allHandles);
    size_t extent = nrc_mrhe_grad_extent(d->mrhe);
    for (int r = 0; r < d->world; r++) {
        NRC_REQUIRE(hs[r].rank == r, "handles are not in rank order");
        NRC_REQUIRE(hs[r].extent == extent, "ranks disagree on the hash-table shape");
        NRC_REQUIRE(hs[r].adam == ((d->mlpAdam ? 1 : 0) | (d->mrheAdam ? 2 : 0) | (d->sharded ? 4 : 0)), "ranks disagree on Adam / the exchange form");
        if (r == d->rank) {
            d->peerGrad[r] = d->mlp->grad; d->peerDelta[r] = d->mrhe->delta; d->peerFlags[r] = d->flags;
            d->peerTables[r] = d->mrhe->tables; d->peerM1[r] = d->mrhe->m1; d->peerM2[r] = d->mrhe->m2;
        } else if (sameProcess) {
            d->peerTables[r] = (float*)(uintptr_t)hs[r].rawTables;
            d->peerM1[r] = (float*)(uintptr_t)hs[r].rawM1; d->peerM2[r] = (float*)(uintptr_t)hs[r].rawM2;
            d->peerGrad[r] = (const float*)(uintptr_t)hs[r].rawGrad;
            d->peerDelta[r] = (const float*)(uintptr_t)hs[r].rawDelta;
            d->peerFlags[r] = (unsigned*)(uintptr_t)hs[r].rawFlags;
        } else {
            NRC_CUDA(cudaIpcOpenMemHandle(&d->opened[r][0], hs[r].grad, cudaIpcMemLazyEnablePeerAccess));
            NRC_CUDA(cudaIpcOpenMemHandle(&d->opened[r][1], hs[r].delta, cudaIpcMemLazyEnablePeerAccess));
            NRC_CUDA(cudaIpcOpenMemHandle(&d->opened[r][2], hs[r].flags, cudaIpcMemLazyEnablePeerAccess));
            d->peerGrad[r] = (const float*)d->opened[r][0];
            d->peerDelta[r] = (const float*)d->opened[r][1];
            d->peerFlags[r] = (unsigned*)d->opened[r][2];
            if (d->sharded) {   // the owner of a slice writes it into this peer's replica
                NRC_CUDA(cudaIpcOpenMemHandle(&d->opened[r][3], hs[r].tables, cudaIpcMemLazyEnablePeerAccess));
                d->peerTables[r] = (float*)d->opened[r][3];
                if (d->mrheAdam) {
                    NRC_CUDA(cudaIpcOpenMemHandle(&d->opened[r][4], hs[r].m1, cudaIpcMemLazyEnablePeerAccess));
                    NRC_CUDA(cudaIpcOpenMemHandle(&d->opened[r][5], hs[r].m2, cudaIpcMemLazyEnablePeerAccess));
                    d->peerM1[r] = (float*)d->opened[r][4]; d->peerM2[r] = (float*)d->opened[r][5];
                }
            }
        }
    }
    d->connected = true;
    d->sameProcess = sameProcess != 0;
    return NRC_OK;
}

static nrc_status dp_launch(nrc_dp* d, int barriers, int clear, bool reduce)
{
    NRC_REQUIRE(d && d->connected, "nrc_dp_connect first");
    nrc_ctx* ctx = d->mlp->ctx;
    DpArgs a;
    memset(&a, 0, sizeof(a));
    for (int r = 0; r < d->world; r++) { a.peerGrad[r] = d->peerGrad[r]; a.peerDelta[r] = d->peerDelta[r]; a.peerFlags[r] = d->peerFlags[r]; }
    a.myFlags = d->flags; a.rank = d->rank; a.world = d->world; a.doStep = reduce ? 1 : 0;
    if (barriers) d->epoch++;
    a.epoch = d->epoch; a.barriers = barriers; a.clear = clear;
    a.params = d->mlp->params; a.grad = d->mlp->grad; a.momentum = d->mlp->momentum; a.paramsT = d->mlp->paramsT;
    a.imgBf16 = d->mlp->paramsBf16;   // kept in step with params whenever it exists, like launch_nrc_step
    a.lr = d->mlp->lr; a.wd = d->mlp->wd; a.beta1 = d->mlp->beta1;
    a.tables = d->mrhe->tables; a.delta = d->mrhe->delta; a.extent = (unsigned)nrc_mrhe_grad_extent(d->mrhe);
    a.mlr = d->mrhe->lr; a.mwd = d->mrhe->wd;
    NRC_REQUIRE(d->mlp->adam == d->mlpAdam && d->mrhe->adam == d->mrheAdam, "Adam was switched after nrc_dp_create: create the exchange again");
    a.sharded = d->sharded ? 1 : 0;
    if (d->sharded) {
        a.nDest = d->world;
        for (int r = 0; r < d->world; r++) { a.destTables[r] = d->peerTables[r]; a.destM1[r] = d->peerM1[r]; a.destM2[r] = d->peerM2[r]; }
    } else {
        a.nDest = 1;
        a.destTables[0] = d->mrhe->tables; a.destM1[0] = d->mrhe->m1; a.destM2[0] = d->mrhe->m2;
    }
    if (reduce && d->mlpAdam) {   // launch_nrc_step's bookkeeping
        d->mlp->steps++;
        a.mlpAdam = 1; a.second = d->mlp->second; a.beta2 = d->mlp->beta2; a.eps = d->mlp->eps;
        a.c1 = 1.0f / (1.0f - powf(d->mlp->beta1, (float)d->mlp->steps)); a.c2 = 1.0f / (1.0f - powf(d->mlp->beta2, (float)d->mlp->steps));
    }
    if (reduce && d->mrheAdam) {  // launch_mrhe_step's
        d->mrhe->steps++;
        a.mrheAdam = 1; a.m1 = d->mrhe->m1; a.m2 = d->mrhe->m2; a.mbeta1 = d->mrhe->beta1; a.mbeta2 = d->mrhe->beta2; a.meps = d->mrhe->eps;
        a.mc1 = 1.0f / (1.0f - powf(d->mrhe->beta1, (float)d->mrhe->steps)); a.mc2 = 1.0f / (1.0f - powf(d->mrhe->beta2, (float)d->mrhe->steps));
    }
    {
        const char* e = getenv("NRC_DP_TIMEOUT_MS");   // default 30 s: longer than any stall of a healthy rank (pool regrowth + re-trace, checkpoint load)
        const long ms = e ? atol(e) : 30000;
        a.timeoutNs = (unsigned long long)(ms > 0 ? ms : 30000) * 1000000ull;
    }
    // latency regime: 16 small blocks; bandwidth regime (sharded): 4 per SM, so that enough NVLink loads and stores are in flight
    unsigned blocks = DP_BLOCKS;
    if (d->sharded) {   // enough CTAs that a thread walks ~8 float4 of its rank's slice: 16 .. 592
        const size_t per4 = (size_t)a.extent / 4 / (size_t)d->world;
        size_t want = per4 / (DP_THREADS * 8) + 1;
        if (const char* e = getenv("NRC_DP_BLOCKS")) want = (size_t)atol(e);
        else if (want > (size_t)ctx->smCount) want = (size_t)ctx->smCount;   // one CTA per SM is the measured optimum (T = 2^22, 2 GPUs:
                                                                              // 0.346 / 0.196 / 0.110 / 0.124 / 0.171 ms with 32 / 64 / 148 / 296 / 592)
        blocks = (unsigned)(want < DP_BLOCKS ? DP_BLOCKS : (want > NRC_DP_MAX_BLOCKS ? NRC_DP_MAX_BLOCKS : want));
    }
    NRC_LAUNCH(ctx, ctx->stream, K_DP_REDUCE_STEP, k_dp_reduce_step<<<blocks, DP_THREADS, 0, ctx->stream>>>(a));
    // the sticky time-out word follows every launch to pinned host memory: the next call / nrc_sync / nrc_readback_wait sees it
    if (barriers) NRC_CUDA(cudaMemcpyAsync(d->hostFail, d->flags + NRC_DP_FLAG_COUNT, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
    return NRC_OK;
}

// One frame's exchange + both optimiser steps.  Multi-process: every rank calls this once per frame, on its own GPU; the
// kernel waits (on the device) for the peers' kernels of the same frame.
extern "C" nrc_status nrc_dp_reduce_step(nrc_dp* d)
{
    NRC_REQUIRE(d && d->connected, "nrc_dp_connect first");
    NRC_REQUIRE(!d->sameProcess || d->world == 1, "same-process replicas: use nrc_dp_reduce_step_phase");
    NRC_TRY(nrc_ctx_check_dp(d->mlp->ctx));
    NRC_TRY(dp_launch(d, d->world > 1, 1, true));
    return launch_mrhe_step_range(d->mrhe, nrc_mrhe_grad_extent(d->mrhe), d->mrhe->ctx->stream);
}

// Same-process replicas (tests): phase 0 = reduce + step without barriers or clearing, for every replica; then phase 1 =
// clear + tail step, for every replica.  The loss accumulator is left per replica in this mode.
extern "C" nrc_status nrc_dp_reduce_step_phase(nrc_dp* d, int phase)
{
    NRC_REQUIRE(d && d->connected, "nrc_dp_connect first");
    NRC_REQUIRE(phase == 0 || phase == 1 || phase == 2, "phase must be 0, 1 or 2");
    if (phase == 2) return dp_launch(d, 1, 1, true);   // test hook: the real kernel with handshakes, on this replica only (time-out test)
    if (phase == 0) return dp_launch(d, 0, 0, true);
    NRC_TRY(dp_launch(d, 0, 2, false));
    return launch_mrhe_step_range(d->mrhe, nrc_mrhe_grad_extent(d->mrhe), d->mrhe->ctx->stream);
}

// *timedOut = 1 if a handshake of any earlier nrc_dp_reduce_step gave up waiting for a peer (waits for the stream).
extern "C" nrc_status nrc_dp_status(nrc_dp* d, int* timedOut)
{
    NRC_REQUIRE(d && timedOut, "null argument");
    unsigned v = 0;
    NRC_CUDA(cudaMemcpyAsync(&v, d->flags + NRC_DP_FLAG_COUNT, sizeof(unsigned), cudaMemcpyDeviceToHost, d->mlp->ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(d->mlp->ctx->stream));
    *timedOut = v != 0u;
    return NRC_OK;
}

extern "C" nrc_status nrc_dp_destroy(nrc_dp* d)
{
    if (!d) return NRC_OK;
    cudaStreamSynchronize(d->mlp->ctx->stream);
    for (int r = 0; r < NRC_DP_MAX_WORLD; r++)
        for (int k = 0; k < 6; k++)
            if (d->opened[r][k]) cudaIpcCloseMemHandle(d->opened[r][k]);
    cudaFree(d->flags);
    {
        std::vector<const volatile unsigned*>& w = d->mlp->ctx->dpFailWords;
        for (size_t i = 0; i < w.size(); i++)
            if (w[i] == d->hostFail) { w.erase(w.begin() + i); break; }
    }
    if (d->hostFail) cudaFreeHost(d->hostFail);
    delete d;
    return NRC_OK;
}
