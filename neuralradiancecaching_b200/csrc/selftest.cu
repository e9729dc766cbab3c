// selftest.cu -- exhaustive device-side checks of the guard-free division / square-root fast paths (strict.cuh) against the IEEE
// routines they shortcut.  Not on any product path: nrc_selftest_strict_math is called by tests/test_gpu_strict_math.py.
#include "nrc_internal.cuh"
#include "strict.cuh"

// [0] grid_coords: (a / extent + 0.5) compared as the range test sees it, ALL 2^32 bit patterns of a, extents 62.5 / 42.5 / 76.5,
//     for div_normal and for the three-instruction div_extent
// [1] sqrt_normal(x) vs __fsqrt_rn(x), every float in [2^-101, FLT_MAX]
// [2] div_normal(1, y) vs __fdiv_rn(1, y), every float in [2^-30, 2^30]
// [3] the two divisions of the HG cos-theta draw: (1-g^2)/den for every den in [0.25, 2), num/(2g) for every |num| <= 2, g = 0.7
__device__ __forceinline__ bool same_coord(float q0, float q1)
{
    float u0 = __fadd_rn(q0, 0.5f), u1 = __fadd_rn(q1, 0.5f);
    bool in0 = (u0 >= -1.0f) && (u0 < 1.0e6f), in1 = (u1 >= -1.0f) && (u1 < 1.0e6f);   // any grid is far below 1e6 cells per axis
    if (in0 != in1) return false;
    return !in0 || __float_as_uint(u0) == __float_as_uint(u1);
}

__global__ void __launch_bounds__(256) k_selftest_strict(unsigned long long* __restrict__ bad)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long t0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long b0 = 0, b1 = 0, b2 = 0, b3 = 0;
    const float ext[3] = {62.5f, 42.5f, 76.5f};
    for (unsigned long long i = t0; i < (1ull << 32); i += stride) {
        float a = __uint_as_float((unsigned)i);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (!same_coord(__fdiv_rn(a, ext[k]), div_normal(a, ext[k]))) b0++;
            if (!same_coord(__fdiv_rn(a, ext[k]), div_extent(a, ext[k], __frcp_rn(ext[k])))) b0++;   // the form grid_coords uses
        }
    }
    for (unsigned long long i = t0 + 0x0D000000ull; i <= 0x7F7FFFFFull; i += stride) {
        float x = __uint_as_float((unsigned)i);
        if (__float_as_uint(__fsqrt_rn(x)) != __float_as_uint(sqrt_normal(x))) b1++;
    }
    for (unsigned long long i = t0 + 0x30800000ull; i <= 0x4E800000ull; i += stride) {
        float y = __uint_as_float((unsigned)i);
        if (__float_as_uint(__fdiv_rn(1.0f, y)) != __float_as_uint(div_normal(1.0f, y))) b2++;
    }
    const float g = 0.7f, gg = __fmul_rn(g, g), num1 = __fsub_rn(1.0f, gg), g2 = __fmul_rn(2.0f, g);
    for (unsigned long long i = t0 + 0x3E800000ull; i < 0x40000000ull; i += stride) {
        float den = __uint_as_float((unsigned)i);
        if (__float_as_uint(__fdiv_rn(num1, den)) != __float_as_uint(div_normal(num1, den))) b3++;
    }
    for (unsigned long long i = t0; i <= 0x40000000ull; i += stride) {      // |num| <= 2, both signs
        for (unsigned sgn = 0; sgn < 2; sgn++) {
            float num = __uint_as_float((unsigned)i | (sgn << 31));
            float c0 = __fdiv_rn(num, g2), c1 = div_normal(num, g2);
            // what the caller does with it: acos(); a quotient below 2^-126 gives pi/2 either way
            if (__float_as_uint(nrc_acos(c0)) != __float_as_uint(nrc_acos(c1))) b3++;
        }
    }
    atomicAdd(bad + 0, b0); atomicAdd(bad + 1, b1); atomicAdd(bad + 2, b2); atomicAdd(bad + 3, b3);
}

extern "C" nrc_status nrc_selftest_strict_math(nrc_ctx* ctx, uint64_t mismatches[4])
{
    NRC_REQUIRE(ctx && mismatches, "null argument");
    unsigned long long* d = nullptr;
    NRC_CUDA(cudaMalloc(&d, 4 * sizeof(unsigned long long)));
    NRC_CUDA(cudaMemsetAsync(d, 0, 4 * sizeof(unsigned long long), ctx->stream));
    NRC_LAUNCH(ctx, ctx->stream, K_MISC, k_selftest_strict<<<ctx->smCount * 16, 256, 0, ctx->stream>>>(d));
    unsigned long long h[4];
    NRC_CUDA(cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    cudaFree(d);
    for (int i = 0; i < 4; i++) mismatches[i] = h[i];
    return NRC_OK;
}
