// mlp_stream.cu -- the 64-wide MLP (Forward, nrc-forward.frag:503-534; ApplyWeights0..5 / ActivateNr1..6, train.comp:547-701)
// as a stand-alone, warp-specialised, software-pipelined tcgen05 kernel: bf16 feature tiles in, RGB out.
//
// Why a second inference kernel next to the fused k_infer_tc: in the fused kernel every tile waits for its 256 hash
// gathers per query before the first MMA, and every layer waits for the previous layer's epilogue, so the tensor pipe
// idles (profiles/r01k: 16 % issue-active, tensor pipe < 10 %).  Here the two halves are separate kernels, as SURVEY 8(d)
// bounds them: k_encode_tiles (mlp_tc.cu) is a pure gather kernel at full occupancy that writes the network inputs as
// bf16 operand-image tiles (16 KB per 128 queries, exactly the shared-memory image the tensor core reads), and
// k_mlp_stream below is bounded by the tensor pipe's round-trip latency:
//
//   one persistent CTA per SM, 5*S + 1 warps, S tiles ("slots") in flight:
//     warps 0..4S-1          epilogue: 4 warps per slot (one per TMEM lane quarter, thread = query row): tcgen05.ld the fp32
//                            accumulator, + bias (packed add.f32x2), ReLU and round to bf16 in one cvt.rn.relu.bf16x2.f32,
//                            tcgen05.st the packed row as the next layer's A operand, arrive on the slot's mbarrier.
//     warps 4S..5S-1 (1 lane) MMA issuer of slot s: wait -> 4 x tcgen05.mma (M=128, N=64, K=16) -> tcgen05.commit.  Layer 0
//                            takes A from the shared-memory stage; layers 1-5 take A **from TMEM** (tcgen05.mma [d], [a], bdesc):
//                            activations never touch shared memory, whose bandwidth (A 4 KB + B 2 KB per 32-cycle MMA) would
//                            otherwise cap the pipe at 2/3.  One issuer PER SLOT: the S dependent chains
//                            (MMA -> commit -> epilogue -> arrive -> MMA ...) run side by side.
//     warp 5S (1 lane)       producer: 1-D bulk copies (cp.async.bulk ... mbarrier::complete_tx) of whole 16 KB tiles into a
//                            ring of NSTAGE shared-memory stages.  NSTAGE is a multiple of S, so a stage is always consumed by
//                            the same slot: its issuer has waited for the stage's previous fill itself, and a parity wait can
//                            never be satisfied by the phase before (no cross-slot ordering assumption).
//   TMEM: slot s owns columns [96 s, 96 s + 64) accumulator D and [96 s + 64, 96 s + 96) bf16 A operand (K = 64 -> 32
//   columns); S = 5 uses 480 of the 512 columns -- TMEM capacity is what bounds the number of chains.  Weights (43 KB bf16
//   image) and biases stay in shared memory for the lifetime of the CTA.
//
// Measured (profiles/r01m_mlp_stream.md): one layer of one tile is a ~780-cycle round trip around 128 cycles of tensor work
// (issue, pipeline, commit -> mbarrier -> wake-up), + ~100-400 cycles of epilogue; five chains give ~250 cycles per tile and layer
// in steady state (~1.0 PFLOP/s) plus ~25 us of launch / fill / drain per launch.  Results are bit-identical to k_infer_tc.
// Work per query 41 344 FLOP, traffic 128 B in + 12 B out.
#include "nrc_internal.cuh"
#include "tc_ptx.cuh"

#define MS_SLOT_COLS 96u

// ---- PTX used only here ----------------------------------------------------------------------------------------------
// D[tmem] (+)= A[tmem] * B[smem]^T : A is 128 lanes x (K/2) 32-bit columns of packed bf16 pairs
__device__ __forceinline__ void umma_bf16_ts(uint32_t dTmem, uint32_t aTmem, uint64_t bDesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
        "}" ::"r"(dTmem), "r"(aTmem), "l"(bDesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* v)
{
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
        "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}

__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// issuer-side wait without the hardware suspend window of try_wait (5 threads per CTA poll; the wake-up is on the chain)
__device__ __forceinline__ void mbar_spin(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}

// ---- shared-memory map ----------------------------------------------------------------------------------------------------
#define MS_SMEM_W 0                                      // forward bf16 weight image, 43 008 B
#define MS_SMEM_BIAS (TC_W_ELEMS * 2)                    // 336 fp32
#define MS_SMEM_BAR (MS_SMEM_BIAS + 336 * 4)             // full[16], sfree[16], accum[8], aready[8]
#define MS_SMEM_SLOT (MS_SMEM_BAR + 48 * 8)
#define MS_SMEM_STAGE (MS_SMEM_SLOT + 128)               // NSTAGE x 16 KB feature tiles (128-byte aligned)
#define MS_SMEM_BYTES(nstage) (MS_SMEM_STAGE + (nstage) * 16384)

template <int S, int NSTAGE, int DBG>
__global__ void __launch_bounds__(160 * S + 32, 1)
    k_mlp_stream(const uint4* __restrict__ featTiles, const uint32_t* __restrict__ countDev, size_t nMax,
                 const __nv_bfloat16* __restrict__ wimg, const float* __restrict__ params, float* __restrict__ rgb)
{
    static_assert(S >= 1 && S <= 5, "a slot is 96 TMEM columns");
    static_assert(NSTAGE >= 2 && NSTAGE <= 11 && NSTAGE % S == 0, "stages of 16 KB beside 43 KB of weights; a stage belongs to one slot");
    extern __shared__ __align__(128) unsigned char smem[];
    float* biasS = reinterpret_cast<float*>(smem + MS_SMEM_BIAS);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + MS_SMEM_BAR);
    uint64_t* sfree = full + 16;
    uint64_t* accum = sfree + 16;
    uint64_t* aready = accum + 8;
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(smem + MS_SMEM_SLOT);

    constexpr int NT = 160 * S + 32;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // ---- one-time setup ----
    {
        const uint4* src = reinterpret_cast<const uint4*>(wimg);
        uint4* dst = reinterpret_cast<uint4*>(smem + MS_SMEM_W);
        for (int i = tid; i < TC_W_ELEMS * 2 / 16; i += NT) dst[i] = __ldg(src + i);
        for (int i = tid; i < 336; i += NT) biasS[i] = (i < 323) ? __ldg(params + NRC_B_OFF(0) + i) : 0.0f;
    }
    if (tid == 0) {
        for (int i = 0; i < NSTAGE; i++) {
            mbar_init(full + i, 1);
            mbar_init(sfree + i, 1);
        }
        for (int i = 0; i < S; i++) {
            mbar_init(accum + i, 1);
            mbar_init(aready + i, 128);   // every epilogue thread arrives (an elected arrive behind __syncwarp measured 5 % slower)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmemSlot, 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmemBase = *tmemSlot;
    const uint32_t wAddr = smem_u32(smem + MS_SMEM_W);

    size_t n = nMax;
    if (countDev) { size_t c = *countDev; n = c < nMax ? c : nMax; }
    const size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    const int nLocal = blockIdx.x < tiles ? (int)((tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;   // tiles b, b+G, ...

    if (warp == 5 * S) {
        // ---- producer ----
        if (lane == 0) {
            for (int j = 0; j < nLocal; j++) {
                const int st = j % NSTAGE;
                if (j >= NSTAGE) mbar_wait(sfree + st, (uint32_t)((j / NSTAGE - 1) & 1));
                mbar_expect_tx(full + st, 16384u);
                const size_t tile = blockIdx.x + (size_t)j * gridDim.x;
                bulk_load(smem + MS_SMEM_STAGE + st * 16384, featTiles + tile * 1024, 16384u, full + st);
            }
        }
    } else if (warp >= 4 * S) {
        // ---- MMA issuer of slot s (one lane): the chain wait -> 4 MMAs -> commit of ONE slot; the S chains run side by side.
        // A single issuer thread for all slots was the bottleneck of the first version (~580 cycles of dependent single-thread
        // instructions per tile and layer around 128 cycles of tensor work); letting a slot's epilogue thread issue behind a named
        // barrier was slower than these dedicated warps (profiles/r01m_mlp_stream.md).
        if (lane == 0) {
            const int s = warp - 4 * S;
            const uint32_t dT = tmemBase + MS_SLOT_COLS * (uint32_t)s;
            // descriptors of the six weight matrices: only the start-address field changes with layer and K step
            const uint64_t bd64 = umma_desc(wAddr, 64 * 16, 128), bd16 = umma_desc(wAddr + 5 * 8192, 16 * 16, 128);
            const uint64_t adBase = umma_desc(smem_u32(smem + MS_SMEM_STAGE), TC_ROWS * 16, 128);
            constexpr uint32_t idesc64 = umma_idesc(TC_ROWS, 64), idesc16 = umma_idesc(TC_ROWS, 16);
            for (int j = s, it = 0; j < nLocal; j += S, it++) {
                const int st = j % NSTAGE;
                if (it > 0) mbar_wait(aready + s, 1u);                        // the previous tile's output has left D[s]
                mbar_wait(full + st, (uint32_t)((j / NSTAGE) & 1));
                tc_fence_after();
                if (!(DBG & 8)) {
                    const uint64_t ad = adBase + (uint64_t)(st * (16384 >> 4));
#pragma unroll
                    for (int ks = 0; ks < 4; ks++)
                        umma_bf16(dT, ad + (uint64_t)(ks * ((2 * TC_ROWS * 16) >> 4)), bd64 + (uint64_t)(ks * ((2 * 64 * 16) >> 4)), idesc64, ks > 0 ? 1u : 0u);
                }
                umma_commit(accum + s);
                umma_commit(sfree + st);                                      // the stage may be refilled
#pragma unroll
                for (int l = 1; l < 6; l++) {
                    if ((DBG & 128) && l < 5) continue;                       // feed probe: layers 0 and 5 only
                    mbar_wait(aready + s, (uint32_t)(((DBG & 128) ? 0 : (l - 1)) & 1));
                    tc_fence_after();
                    if (!(DBG & 8)) {
#pragma unroll
                        for (int ks = 0; ks < 4; ks++) {
                            const uint64_t bd = (l < 5) ? bd64 + (uint64_t)((l * 8192 + ks * 2 * 64 * 16) >> 4) : bd16 + (uint64_t)((ks * 2 * 16 * 16) >> 4);
                            umma_bf16_ts(dT, dT + 64u + 8u * (uint32_t)ks, bd, (l < 5) ? idesc64 : idesc16, ks > 0 ? 1u : 0u);
                        }
                    }
                    umma_commit(accum + s);
                }
            }
        }
    } else {
        // ---- epilogue: slot = warp / 4, TMEM lane quarter = warp % 4, thread = query row ----
        const int s = warp >> 2, quad = warp & 3;
        const int row = quad * 32 + lane;
        const uint32_t dT = tmemBase + MS_SLOT_COLS * (uint32_t)s + ((uint32_t)(quad * 32) << 16);
        for (int j = s; j < nLocal; j += S) {
            const size_t q = (blockIdx.x + (size_t)j * gridDim.x) * TC_ROWS + row;
            uint32_t ph = 0;
#pragma unroll 1
            for (int l = 0; l < 6; l++) {
                if ((DBG & 128) && l >= 1 && l < 5) continue;
                mbar_wait(accum + s, ph);
                ph ^= 1u;
                tc_fence_after();
                if (l < 5) {
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        // one half (32 accumulator columns) of the row: + bias, ReLU, bf16 pairs -> 16 columns of the A operand
                        uint32_t v[32], w[16];
                        if (DBG & 1) {
#pragma unroll
                            for (int i = 0; i < 32; i++) v[i] = (uint32_t)(l + i + h);
                        } else {
                            tmem_ld32(dT + (uint32_t)(h * 32), v);
                            tmem_ld_wait();
                        }
                        const float4* b = reinterpret_cast<const float4*>(biasS + l * 64 + h * 32);
#pragma unroll
                        for (int c = 0; c < 8; c++) {
                            if (DBG & 2) { w[2 * c] = v[4 * c]; w[2 * c + 1] = v[4 * c + 2]; continue; }
                            float4 bb = b[c];
                            float x0, x1, x2, x3;
                            add2_bits(x0, x1, v[4 * c], v[4 * c + 1], bb.x, bb.y);
                            add2_bits(x2, x3, v[4 * c + 2], v[4 * c + 3], bb.z, bb.w);
                            w[2 * c] = pack_relu_bf16(x0, x1);
                            w[2 * c + 1] = pack_relu_bf16(x2, x3);
                        }
                        if (DBG & 4) {
                            uint32_t x = 0;
#pragma unroll
                            for (int i = 0; i < 16; i++) x ^= w[i];
                            if (x == 0x12345678u) rgb[0] = 1.0f;   // keeps the arithmetic alive
                        } else {
                            tmem_st16(dT + 64u + (uint32_t)(h * 16), w);
                        }
                    }
                    if (!(DBG & 4)) tmem_st_wait();
                } else {
                    uint32_t v[4];
                    tmem_ld4(dT, v);
                    tmem_ld_wait();
                    if (q < n) {
#pragma unroll
                        for (int c = 0; c < 3; c++) rgb[3 * q + c] = fmaxf(__uint_as_float(v[c]) + biasS[320 + c], 0.0f);
                    }
                }
                tc_fence_before();
                mbar_arrive(aready + s);
            }
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) tmem_dealloc(tmemBase, 512);
}

// ---- host launcher ---------------------------------------------------------------------------------------------------------
nrc_status tc_prepare(nrc_mlp* mlp, cudaStream_t st);   // mlp_tc.cu

static int ms_slots()   // NRC_MLP_SLOTS: tiles in flight per CTA (read at every launch so that a sweep can change it)
{
    const char* e = getenv("NRC_MLP_SLOTS");
    int s = e ? atoi(e) : 5;
    return s < 1 ? 1 : (s > 5 ? 5 : s);
}

template <int S, int NSTAGE, int DBG>
static nrc_status ms_launch(nrc_mlp* mlp, const void* featTiles, const uint32_t* countDev, size_t n, float* rgb, cudaStream_t st)
{
    static bool attr[64] = {};   // per device: function attributes belong to the device's context
    if (first_on_device(mlp->ctx, attr)) {
        NRC_CUDA(cudaFuncSetAttribute(k_mlp_stream<S, NSTAGE, DBG>, cudaFuncAttributeMaxDynamicSharedMemorySize, MS_SMEM_BYTES(NSTAGE)));
    }
    size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    size_t cap = (size_t)mlp->ctx->smCount;   // persistent: one CTA per SM owns all 512 TMEM columns
    unsigned blocks = (unsigned)(tiles < cap ? tiles : cap);
    NRC_LAUNCH(mlp->ctx, st, K_MLP_STREAM,
               k_mlp_stream<S, NSTAGE, DBG><<<blocks, 160 * S + 32, MS_SMEM_BYTES(NSTAGE), st>>>(
                   reinterpret_cast<const uint4*>(featTiles), countDev, n, reinterpret_cast<const __nv_bfloat16*>(mlp->paramsBf16), mlp->params, rgb));
    return NRC_OK;
}

nrc_status launch_mlp_stream(nrc_mlp* mlp, const void* featTiles, const uint32_t* countDev, size_t n, float* rgb, cudaStream_t st)
{
    if (n == 0) return NRC_OK;
    NRC_TRY(tc_prepare(mlp, st));
#ifdef NRC_MLP_PROBES   // bottleneck probes (WRONG results by construction, scripts/mlp_stream_probes.py): NRC_MLP_DBG bit 0 no tcgen05.ld,
                        // bit 1 no arithmetic, bit 2 no tcgen05.st, bit 3 no MMA, bit 7 layers 0 and 5 only; x00: x slots
    if (const char* e = getenv("NRC_MLP_DBG")) {
        switch (atoi(e)) {
        case 7: return ms_launch<5, 10, 7>(mlp, featTiles, countDev, n, rgb, st);
        case 8: return ms_launch<5, 10, 8>(mlp, featTiles, countDev, n, rgb, st);
        case 128: return ms_launch<5, 10, 128>(mlp, featTiles, countDev, n, rgb, st);
        case 107: return ms_launch<1, 10, 7>(mlp, featTiles, countDev, n, rgb, st);
        case 108: return ms_launch<1, 10, 8>(mlp, featTiles, countDev, n, rgb, st);
        default: break;
        }
    }
#endif
    switch (ms_slots()) {
    case 1: return ms_launch<1, 10, 0>(mlp, featTiles, countDev, n, rgb, st);
    case 2: return ms_launch<2, 10, 0>(mlp, featTiles, countDev, n, rgb, st);
    case 3: return ms_launch<3, 9, 0>(mlp, featTiles, countDev, n, rgb, st);
    case 4: return ms_launch<4, 8, 0>(mlp, featTiles, countDev, n, rgb, st);
    default: return ms_launch<5, 10, 0>(mlp, featTiles, countDev, n, rgb, st);
    }
}
