// mrhe_scatter.cuh -- BackpropMrhe (nrc-train.comp:947-983): the hash-table gradient scatter of both train kernels.
//
// The reference adds one float atomic per corner, per feature and per sample into the delta table -- 256 atomics per sample, and in
// the shipped arithmetic all 16 levels of a sample land in ONE 2*T-float slice with the same value for both features.
// At 2^18 samples that is 33.5 M contended global atomics onto 16 384 addresses, which held the tensor-core train kernel at 3 % of the
// tensor roofline (round 1).  Here the scatter is the LAST phase of the train kernel, after all its tiles:
//
//   phase 1 (per tile, in the train kernels): the 32 hash-feature gradients dL/dx0[0..31] of every sample go to a global scratch row
//            (128 B per sample, L2-resident); dropped samples (NaN guard, rows past n) write zeros.
//   phase 2 (here): the CTA walks its tiles again, recomputes the corner indices and weights (integer hashes + three subtractions;
//            the same strict functions as the encoder, so the addresses are bit-identical), and accumulates
//              * parity preset, T <= NRC_HIST_MAX_T: into a PRIVATE histogram of T floats in the CTA's shared memory (64 KiB at
//                T = 2^14; both features receive the same value, A9, so one float per entry is enough; the activations that lived
//                there are dead by now).  Shared-memory float adds, no global atomic at all.  The histogram leaves the CTA as one
//                coalesced store into its own partial row; k_reduce_partials adds the rows in fixed CTA order into the delta table
//                (the same fixed-order reduction the MLP gradients get).
//              * otherwise (tables too large for shared memory: T = 2^22 is 16 MiB per level; or the opt-in fixed preset, whose
//                gradients go to 16 per-level tables): global red.add.v2.f32, warp-aggregated on the coarse levels -- lanes of a warp
//                that hit the same entry (__match_any_sync on the hashed index) are summed in lane order and the lowest lane issues
//                one atomic.  On the fine levels (more cells than a warp can plausibly share) the match would be pure overhead.
//
// Values: every contribution is ((-g * ((xf * yf) * zf)) * scale), the reference's expression (A11: un-swapped axes, unbounded
// factors); only the order of the float additions differs from the oracle's loop, as with any atomic implementation.
#pragma once
#include "nrc_internal.cuh"
#include "strict.cuh"

#define NRC_HIST_MAX_T 32768u      // 128 KiB of shared memory
#define NRC_AGG_MAX_RES 64u        // warp aggregation for levels with at most 64^3 cells

// sum of `v` over the lanes in `peers` (all of which call this with the same mask), in ascending lane order
__device__ __forceinline__ float peers_sum(unsigned peers, float v)
{
    float s = 0.0f;
    for (unsigned mm = peers; mm; mm &= mm - 1u) s += __shfl_sync(peers, v, __ffs((int)mm) - 1);
    return s;
}

// ROWS samples per tile (64: fp32 kernel, 128: tcgen05 kernel), THREADS threads: thread (row, part) owns 16 / (THREADS / ROWS) levels of its row.
// gAll[q][32]: dL/dx0 of the 32 hash features of sample q (zeros for dropped samples).  hist: shared memory, T floats (or nullptr:
// global path).  histOut: this CTA's partial row in global memory (T floats), written when hist != nullptr.
template <int ROWS, int THREADS = 256>
__device__ __forceinline__ void mrhe_scatter_tiles(const float* __restrict__ pos, const float* gAll, size_t n, const MrheDev& m,
                                                   float scale, float* hist, float* __restrict__ histOut)
{
    constexpr int PARTS = THREADS / ROWS;
    constexpr int LEVELS = 16 / PARTS;
    const int tid = threadIdx.x;
    const int row = tid % ROWS, part = tid / ROWS;
    const size_t tiles = (n + ROWS - 1) / ROWS;
    const f3 sky = mk3(m.skySize[0], m.skySize[1], m.skySize[2]);
    if (hist) {
        for (uint32_t i = tid; i < m.T / 4; i += THREADS) reinterpret_cast<float4*>(hist)[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        __syncthreads();
    }
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t q = tile * ROWS + row;
        const bool live = q < n;
        f3 p = mk3(0.0f, 0.0f, 0.0f);
        float g[2 * LEVELS];
#pragma unroll
        for (int i = 0; i < 2 * LEVELS; i++) g[i] = 0.0f;
        if (live) {
            p = mk3(__ldg(pos + 3 * q), __ldg(pos + 3 * q + 1), __ldg(pos + 3 * q + 2));
            const float4* gs = reinterpret_cast<const float4*>(gAll + q * 32 + 2 * LEVELS * part);
#pragma unroll
            for (int i = 0; i < LEVELS / 2; i++) {
                float4 v = __ldcg(gs + i);   // written by this CTA in phase 1: read at L2, never through the read-only path
                g[4 * i] = v.x; g[4 * i + 1] = v.y; g[4 * i + 2] = v.z; g[4 * i + 3] = v.w;
            }
        }
        const f3 np = mrhe_norm_pos(p, sky);
        if (hist) {
            bool any = false;
#pragma unroll
            for (int i = 0; i < LEVELS; i++) any |= (g[2 * i] != 0.0f);
            if (!any) continue;   // dropped sample / zero error: nothing to add
#pragma unroll 1
            for (int li = 0; li < LEVELS; li++) {
                const uint32_t level = LEVELS * part + li;
                if (level >= m.L) break;
                const float ng = -g[2 * li];   // feature 0's error for both features (A9)
                if (ng == 0.0f) continue;
                uint32_t idx[8];
                f3 lf;
                mrhe_level_indices(p, np, m.res[level], m.T, idx, lf, false);
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    const float xf = (c & 4) ? lf.x : (1.0f - lf.x);
                    const float yf = (c & 2) ? lf.y : (1.0f - lf.y);
                    const float zf = (c & 1) ? lf.z : (1.0f - lf.z);
                    atomicAdd(hist + idx[c], (ng * ((xf * yf) * zf)) * scale);   // no level offset (A10)
                }
            }
        } else {
            // global path: the whole warp stays converged (dead lanes add zeros) so that the match below is legal
            float2* dtab = reinterpret_cast<float2*>(m.delta);
            const bool fixed = m.fixed != 0;
#pragma unroll 1
            for (int li = 0; li < LEVELS; li++) {
                const uint32_t level = LEVELS * part + li;
                if (level >= m.L) break;
                const float ng = -g[2 * li];
                const float ng1 = fixed ? -g[2 * li + 1] : ng;   // fixed preset: feature 1 gets its own error
                uint32_t idx[8];
                f3 lf;
                mrhe_level_indices(p, np, m.res[level], m.T, idx, lf, fixed);
                float2* dl = dtab + (fixed ? (size_t)level * m.T : 0);   // fixed preset: its level's table
                const bool aggregate = m.res[level] <= NRC_AGG_MAX_RES;
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    const float xf = (c & 4) ? lf.x : (1.0f - lf.x);
                    const float yf = (c & 2) ? lf.y : (1.0f - lf.y);
                    const float zf = (c & 1) ? lf.z : (1.0f - lf.z);
                    const float wgt = (xf * yf) * zf;
                    float d0 = (ng * wgt) * scale, d1 = (ng1 * wgt) * scale;
                    if (aggregate) {
                        const unsigned peers = __match_any_sync(0xffffffffu, idx[c]);
                        d0 = peers_sum(peers, d0);
                        d1 = peers_sum(peers, d1);
                        if ((int)(tid & 31) != __ffs((int)peers) - 1) continue;
                    }
                    if (d0 != 0.0f || d1 != 0.0f) atomicAdd(dl + idx[c], make_float2(d0, d1));
                }
            }
        }
    }
    if (hist) {
        __syncthreads();
        for (uint32_t i = tid; i < m.T / 4; i += THREADS) reinterpret_cast<float4*>(histOut)[i] = reinterpret_cast<const float4*>(hist)[i];
    }
}
