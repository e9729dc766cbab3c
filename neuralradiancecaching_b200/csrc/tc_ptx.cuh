// tc_ptx.cuh -- tcgen05 / TMEM / mbarrier PTX wrappers and operand-image constants shared by the tensor-core kernels
// (mlp_tc.cu: fused encode + MLP; mlp_stream.cu: the pipelined MLP-only kernel).
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

#define TC_ROWS 128
#define TC_THREADS 256
#define TC_TMEM_COLS 64

// ---- bf16 weight images ----------------------------------------------------------------------------------------------
// forward image (B operand of x*W^T: rows n = out, K = in), elements:
//   layer l < 5 at l*4096: (k/8)*512 + n*8 + k%8 ; layer 5 (3 rows padded to 16) at 20480: (k/8)*128 + n*8 + k%8
// transposed image at TC_W_ELEMS (B operand of the back-data GEMM e*W: rows n = in, K = out):
//   layer l < 5 at l*4096: (k/8)*512 + n*8 + k%8 with k = out ; layer 5 at 20480: K = 16 (3 real outputs), same formula
#define TC_W_ELEMS (5 * 4096 + 1024)
// The training kernel's FORWARD pass runs in split precision (bf16x3: w = hi + mid + lo, each a bf16; six partial products per K step
// reproduce the fp32 product to ~2^-24): two more forward images, same layout as the first, hold the mid and lo parts.
#define TC_IMG_T TC_W_ELEMS              // transposed image
#define TC_IMG_MID (2 * TC_W_ELEMS)      // forward image, mid part
#define TC_IMG_LO (3 * TC_W_ELEMS)       // forward image, lo part
#define TC_IMG_ELEMS (4 * TC_W_ELEMS)

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

__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

// whole-tile 1-D bulk copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_load(void* dstSmem, const void* srcGlobal, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dstSmem)),
                 "l"(srcGlobal), "r"(bytes), "r"(smem_u32(bar))
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

// operand majorness bits of the instruction descriptor: set = MN-major (the M / N index is the contiguous one), clear = K-major
#define UMMA_A_MN (1u << 15)
#define UMMA_B_MN (1u << 16)

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

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
// COLS = 32 or 16 consecutive columns
template <int COLS>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t* v)
{
    if (COLS == 32) tmem_ld32(taddr, v); else tmem_ld16(taddr, v);
}

__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t* v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                 : "r"(taddr)
                 : "memory");
}

__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// relu(hi), relu(lo) rounded to nearest-even bf16, hi in the upper half
__device__ __forceinline__ uint32_t pack_relu_bf16(float lo, float hi)
{
    uint32_t d;
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}

// packed fp32 pair add (Blackwell FADD2): {a.x + b.x, a.y + b.y}
__device__ __forceinline__ void add2_bits(float& x0, float& x1, uint32_t a0, uint32_t a1, float b0, float b1)
{
    asm("{\n\t"
        ".reg .b64 ra, rb, rc;\n\t"
        "mov.b64 ra, {%2, %3};\n\t"
        "mov.b64 rb, {%4, %5};\n\t"
        "add.rn.f32x2 rc, ra, rb;\n\t"
        "mov.b64 {%0, %1}, rc;\n\t"
        "}"
        : "=f"(x0), "=f"(x1)
        : "r"(a0), "r"(a1), "f"(b0), "f"(b1));
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// x = hi + mid + lo with three bf16 parts (round to nearest each; the residuals are exact in fp32)
__device__ __forceinline__ void split_bf16x3(float x, unsigned short& hi, unsigned short& mid, unsigned short& lo)
{
    hi = __bfloat16_as_ushort(__float2bfloat16_rn(x));
    const float r1 = x - __uint_as_float((uint32_t)hi << 16);
    mid = __bfloat16_as_ushort(__float2bfloat16_rn(r1));
    const float r2 = r1 - __uint_as_float((uint32_t)mid << 16);
    lo = __bfloat16_as_ushort(__float2bfloat16_rn(r2));
}

// the same for a pair, packed like pack_bf16 (first value in the low half)
__device__ __forceinline__ void split2_bf16x3(float a, float b, uint32_t& hi, uint32_t& mid, uint32_t& lo)
{
    hi = pack_bf16(a, b);
    const float ra = a - __uint_as_float(hi << 16), rb = b - __uint_as_float(hi & 0xffff0000u);
    mid = pack_bf16(ra, rb);
    const float sa = ra - __uint_as_float(mid << 16), sb = rb - __uint_as_float(mid & 0xffff0000u);
    lo = pack_bf16(sa, sb);
}

// One MLP weight (packed arena index i < 20 672, value w) into all four operand images (layouts above): the forward image and its
// mid / lo parts, and the transposed image.  Called by k_pack_bf16 and by every optimiser-step kernel, so the images never go stale.
__device__ __forceinline__ void tc_store_weight(unsigned short* __restrict__ img, int i, float w)
{
    const int l = i >> 12, rem = i & 4095, o = rem >> 6, k = rem & 63;
    const int dst = (l < 5) ? (l * 4096 + (k >> 3) * 512 + o * 8 + (k & 7)) : (5 * 4096 + (k >> 3) * 128 + o * 8 + (k & 7));
    unsigned short hi, mid, lo;
    split_bf16x3(w, hi, mid, lo);
    img[dst] = hi;
    img[TC_IMG_MID + dst] = mid;
    img[TC_IMG_LO + dst] = lo;
    img[TC_IMG_T + l * 4096 + (o >> 3) * 512 + k * 8 + (o & 7)] = hi;   // rows = in, K = out; layer 5 has K = 16
}

