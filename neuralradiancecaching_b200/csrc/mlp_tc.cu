// mlp_tc.cu -- fused MRHE/OneBlob encode + 64-wide MLP on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// SURVEY.md 8a rows a9-a11 (EncodeRay + Forward, nrc-forward.frag:503-534) as one kernel per query batch:
//
//   CTA tile = 128 queries = the 128 TMEM lanes.  256 threads: thread (row, half) encodes half of the 64 network
//   inputs of its row (8 MRHE levels + 16 OneBlob bins) straight into the A operand in shared memory (bf16).
//   Each hidden layer is D[128x64] = A[128x64] * W_l^T: four tcgen05.mma.kind::f16 (M=128, N=64, K=16) issued by ONE
//   thread, fp32 accumulators in TMEM, completion signalled through tcgen05.commit -> mbarrier.  The epilogue reads
//   the accumulator with tcgen05.ld (warp w owns TMEM lanes 32*(w%4).., column half w/4), adds the fp32 bias, applies
//   ReLU, rounds to bf16 and writes the next layer's A operand back to the same shared-memory tile.  All six weight
//   matrices stay resident in shared memory (43 KB bf16) for the lifetime of the CTA; ~61 KB of shared memory and 64
//   TMEM columns per CTA let three CTAs share an SM, so one CTA's gathers/epilogue overlap another's MMAs.
//
// Operand layout (both operands K-major, no swizzle = "interleave" canonical layout of the UMMA shared-memory
// descriptor): a core matrix is 8 rows x 16 bytes (8 bf16 along K), stored as 128 contiguous bytes;
//     byte(row, k) = (k / 8) * (ROWS * 16) + row * 16 + (k % 8) * 2
// so LBO (core matrices adjacent in K) = ROWS*16 and SBO (adjacent 8-row groups) = 128.  A thread writing the 16-byte
// chunk of its own row gives 512 contiguous bytes per warp: conflict-free, no swizzle arithmetic.
//
// Precision: operands are bf16 (8-bit mantissa), accumulation fp32, biases fp32.  Against the fp32 oracle the stated
// bound is 3e-2 of the largest output magnitude (tests/test_gpu_tc.py); the fp32 CUDA-core kernels (mlp_fp32.cu) stay
// the 1e-3 parity path.
#include <cuda_bf16.h>
#include "nrc_internal.cuh"
#include "strict.cuh"

#define TC_ROWS 128
#define TC_THREADS 256
#define TC_TMEM_COLS 64

// ---- bf16 weight image -----------------------------------------------------------------------------------------------
// elements: layer l < 5 at l*4096: (k/8)*512 + n*8 + k%8 ; layer 5 (3 rows padded to 16) at 20480: (k/8)*128 + n*8 + k%8
#define TC_W_ELEMS (5 * 4096 + 1024)

__global__ void __launch_bounds__(256) k_pack_bf16(const float* __restrict__ params, __nv_bfloat16* __restrict__ img)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= TC_W_ELEMS) return;
    float v = 0.0f;
    if (i < 5 * 4096) {
        int l = i >> 12, rem = i & 4095, kc = rem >> 9, n = (rem >> 3) & 63, kk = rem & 7;
        v = params[l * 4096 + n * 64 + kc * 8 + kk];
    } else {
        int rem = i - 5 * 4096, kc = rem >> 7, n = (rem >> 3) & 15, kk = rem & 7;
        if (n < 3) v = params[NRC_W_OFF(5) + n * 64 + kc * 8 + kk];
    }
    img[i] = __float2bfloat16_rn(v);
}

// ---- PTX wrappers ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols)   // one full warp
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols)   // the same warp
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (the tensor core reads operands through it)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, kind::f16 (bf16 in, fp32 accumulate), one CTA
__device__ __forceinline__ void umma_bf16(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(dTmem), "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// all MMAs issued so far by this thread -> one arrive on the mbarrier when they have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// shared-memory matrix descriptor, K-major, no swizzle (layout_type 0), version 1 (Blackwell)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smemAddr, uint32_t lboBytes, uint32_t sboBytes)
{
    uint64_t d = 0;
    d |= (uint64_t)((smemAddr >> 4) & 0x3fffu);
    d |= (uint64_t)((lboBytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sboBytes >> 4) & 0x3fffu) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}

// instruction descriptor: D fp32 (bit 4), A bf16 (bit 7), B bf16 (bit 10), both K-major, N>>3 at 17, M>>4 at 24
__device__ __forceinline__ constexpr uint32_t umma_idesc(uint32_t M, uint32_t N)
{
    return (1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// 32 lanes x 32 consecutive fp32 columns: thread = its lane's row
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t* v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                 : "r"(taddr)
                 : "memory");
}

__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// ---- shared-memory map ----------------------------------------------------------------------------------------------------
#define TC_SMEM_W 0                               // bf16 weight image, 43 008 B
#define TC_SMEM_A (TC_W_ELEMS * 2)                // A tile: 8 K-chunks x 128 rows x 16 B = 16 384 B
#define TC_SMEM_BIAS (TC_SMEM_A + 16384)          // 323 fp32 biases (336 slots)
#define TC_SMEM_BAR (TC_SMEM_BIAS + 336 * 4)      // mbarrier (8 B) + TMEM base slot (4 B)
#define TC_SMEM_BYTES (TC_SMEM_BAR + 16)

// Encode half of row `r`'s network inputs into the A tile.  half 0: MRHE levels 0-7 + OneBlob theta; half 1: levels
// 8-15 + OneBlob phi.  Index arithmetic and blends are the strict ones (bit-identical hash indices); only the final
// rounding to bf16 differs from the fp32 path.
__device__ __forceinline__ void tc_encode_half(unsigned char* aTile, int r, int half, bool live, f3 p, f3 d, const MrheDev& m)
{
    uint4* chunk = reinterpret_cast<uint4*>(aTile);   // chunk(kc, row) = chunk[kc * 128 + row]
    f3 sky = mk3(m.skySize[0], m.skySize[1], m.skySize[2]);
    f3 np = mrhe_norm_pos(p, sky);
#pragma unroll
    for (int g = 0; g < 2; g++) {
        float o[8];
#pragma unroll
        for (int li = 0; li < 4; li++) {
            uint32_t level = 8 * half + 4 * g + li;
            float o0 = 0.0f, o1 = 0.0f;
            if (live && level < m.L) {
                uint32_t idx[8];
                f3 lerp;
                mrhe_level_indices(p, np, m.res[level], m.T, idx, lerp);
                const float2* tab = reinterpret_cast<const float2*>(m.tables) + (size_t)level * m.T;
                float f0[8], f1[8];
#pragma unroll
                for (int c = 0; c < 8; c++) {
                    float2 v = __ldg(tab + idx[c]);
                    f0[c] = v.x;
                    f1[c] = v.y;
                }
                o0 = mrhe_blend(f0, lerp);
                o1 = mrhe_blend(f1, lerp);
            }
            o[2 * li] = o0;
            o[2 * li + 1] = o1;
        }
        chunk[(2 * half + g) * TC_ROWS + r] = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
    }
    float ang;
    if (half == 0) {
        ang = NRC_ADD(NRC_DIV(nrc_atan2(d.z, d.x), NRC_PI), 0.5f);
    } else {
        float lenxz = NRC_SQRT(NRC_ADD(NRC_MUL(d.x, d.x), NRC_MUL(d.z, d.z)));
        ang = NRC_ADD(NRC_DIV(nrc_atan2(lenxz, d.y), NRC_PI), 0.5f);
    }
    const float sigma = 0.25f;
    const float term1 = NRC_DIV(1.0f, NRC_MUL(sigma, NRC_SQRT(NRC_MUL(2.0f, NRC_PI))));
#pragma unroll
    for (int g = 0; g < 2; g++) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            float fi = (float)(8 * g + i);
            float t2 = NRC_DIV(NRC_SUB(fi, ang), sigma);
            o[i] = live ? NRC_MUL(term1, nrc_exp(NRC_MUL(NRC_MUL(-0.5f, t2), t2))) : 0.0f;
        }
        chunk[(4 + 2 * half + g) * TC_ROWS + r] = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
    }
}

__global__ void __launch_bounds__(TC_THREADS, 3) k_infer_tc(const float* __restrict__ pos, const float* __restrict__ dir,
                                                            const uint32_t* __restrict__ countDev, size_t nMax,
                                                            const __nv_bfloat16* __restrict__ wimg, const float* __restrict__ params, MrheDev m,
                                                            float* __restrict__ rgb)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* aTile = smem + TC_SMEM_A;
    float* biasS = reinterpret_cast<float*>(smem + TC_SMEM_BIAS);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + TC_SMEM_BAR);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(smem + TC_SMEM_BAR + 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;

    // ---- one-time setup: weights + biases -> smem, mbarrier, TMEM ----
    {
        const uint4* src = reinterpret_cast<const uint4*>(wimg);
        uint4* dst = reinterpret_cast<uint4*>(smem + TC_SMEM_W);
        for (int i = tid; i < TC_W_ELEMS * 2 / 16; i += TC_THREADS) dst[i] = __ldg(src + i);
        for (int i = tid; i < 336; i += TC_THREADS) biasS[i] = (i < 323) ? __ldg(params + NRC_B_OFF(0) + i) : 0.0f;
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmemSlot, TC_TMEM_COLS);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmemBase = *tmemSlot;
    const uint32_t aAddr = smem_u32(aTile), wAddr = smem_u32(smem + TC_SMEM_W);
    uint32_t phase = 0;

    size_t n = nMax;
    if (countDev) { size_t c = *countDev; n = c < nMax ? c : nMax; }
    const size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t q = tile * TC_ROWS + row;
        const bool live = q < n;
        f3 p = mk3(0.0f, 0.0f, 0.0f), d = mk3(0.0f, 0.0f, 1.0f);
        if (live) {
            p = mk3(__ldg(pos + 3 * q), __ldg(pos + 3 * q + 1), __ldg(pos + 3 * q + 2));
            d = mk3(__ldg(dir + 3 * q), __ldg(dir + 3 * q + 1), __ldg(dir + 3 * q + 2));
        }
        tc_encode_half(aTile, row, half, live, p, d, m);
        fence_async_smem();
        tc_fence_before();
        __syncthreads();
#pragma unroll 1
        for (int l = 0; l < 6; l++) {
            if (tid == 0) {
                tc_fence_after();
                const uint32_t nOut = (l < 5) ? 64u : 16u;
                const uint32_t idesc = umma_idesc(TC_ROWS, nOut);
                const uint32_t wBase = wAddr + (uint32_t)l * 8192u;          // layer 5 starts at 5*8192 = 40960
                const uint32_t wLbo = nOut * 16u;                             // 64 rows -> 1024 B, 16 rows -> 256 B
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
                    uint64_t ad = umma_desc(aAddr + ks * 2 * (TC_ROWS * 16), TC_ROWS * 16, 128);
                    uint64_t bd = umma_desc(wBase + ks * 2 * wLbo, wLbo, 128);
                    umma_bf16(tmemBase, ad, bd, idesc, ks > 0 ? 1u : 0u);
                }
                umma_commit(bar);
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            tc_fence_after();
            if (l < 5) {
                uint32_t v[32];
                tmem_ld32(tmemBase + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 32), v);
                tmem_ld_wait();
                const float* b = biasS + l * 64 + half * 32;
                uint4* chunk = reinterpret_cast<uint4*>(aTile);
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    float x[8];
#pragma unroll
                    for (int j = 0; j < 8; j++) x[j] = fmaxf(__uint_as_float(v[8 * c + j]) + b[8 * c + j], 0.0f);
                    chunk[(4 * half + c) * TC_ROWS + row] =
                        make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7]));
                }
                fence_async_smem();
            } else if (half == 0) {
                uint32_t v[4];
                tmem_ld4(tmemBase + ((uint32_t)(quad * 32) << 16), v);
                tmem_ld_wait();
                if (live) {
#pragma unroll
                    for (int j = 0; j < 3; j++) rgb[3 * q + j] = fmaxf(__uint_as_float(v[j]) + biasS[320 + j], 0.0f);
                }
            }
            tc_fence_before();
            __syncthreads();
        }
    }
    tc_fence_after();
    if (warp == 0) tmem_dealloc(tmemBase, TC_TMEM_COLS);
}

// ---- host launchers -----------------------------------------------------------------------------------------------------------
static nrc_status tc_prepare(nrc_mlp* mlp, cudaStream_t st)
{
    if (!mlp->paramsBf16) NRC_CUDA(cudaMalloc(&mlp->paramsBf16, TC_W_ELEMS * sizeof(uint16_t)));
    if (mlp->dirtyBf16) {
        NRC_LAUNCH(mlp->ctx, st, K_PACK_BF16,
                   k_pack_bf16<<<(TC_W_ELEMS + 255) / 256, 256, 0, st>>>(mlp->params, reinterpret_cast<__nv_bfloat16*>(mlp->paramsBf16)));
        mlp->dirtyBf16 = false;
    }
    return NRC_OK;
}

nrc_status launch_infer_tc(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, float* rgb,
                           cudaStream_t st)
{
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    NRC_TRY(tc_prepare(mlp, st));
    static bool attr = false;
    if (!attr) {
        NRC_CUDA(cudaFuncSetAttribute(k_infer_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
        attr = true;
    }
    size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    size_t cap = (size_t)mlp->ctx->smCount * 3;   // persistent: 3 CTAs per SM, grid-stride over the tiles
    unsigned blocks = (unsigned)(tiles < cap ? tiles : cap);
    NRC_LAUNCH(mlp->ctx, st, K_INFER_TC,
               k_infer_tc<<<blocks, TC_THREADS, TC_SMEM_BYTES, st>>>(pos, dir, countDev, n, reinterpret_cast<const __nv_bfloat16*>(mlp->paramsBf16),
                                                                     mlp->params, make_mrhe_dev(m), rgb));
    return NRC_OK;
}
