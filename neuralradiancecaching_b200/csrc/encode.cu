// encode.cu -- stand-alone MRHE / OneBlob encoders.
//
// MRHE forward = EncodePosMrhe (data/shader/nrc-train/nrc-train.comp:266-339): per query 16 levels x 8
// hashed corners x float2.  One thread per (query, level): a warp covers 2 queries, its 32 float2 results are
// one contiguous 256-byte row pair (coalesced store), its 8 gathers per thread are independent 8-byte
// read-only loads (ld.global.nc.v2) that hit L2/L1 -- the tables (2 MiB at T=2^14) never leave L2.
// Algorithmic bytes: 12 (pos) + 1024 (table reads) + 128 (features) = 1164 B/query; bounded by L2 gather
// rate, reported against HBM peak as SURVEY.md 8d asks.
#include "nrc_internal.cuh"
#include "strict.cuh"

__global__ void __launch_bounds__(256) k_encode_pos(const float* __restrict__ pos, size_t n, MrheDev m, float* __restrict__ feat,
                                                    uint32_t* __restrict__ idxOut)
{
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t q = t >> 4;
    uint32_t level = (uint32_t)(t & 15);
    if (q >= n) return;
    f3 p = mk3(__ldg(pos + 3 * q), __ldg(pos + 3 * q + 1), __ldg(pos + 3 * q + 2));
    float2 out = make_float2(0.0f, 0.0f);
    uint32_t idx[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (level < m.L) {
        f3 sky = mk3(m.skySize[0], m.skySize[1], m.skySize[2]);
        f3 np = mrhe_norm_pos(p, sky);
        f3 lerp;
        mrhe_level_indices(p, np, m.res[level], m.T, idx, lerp, m.fixed != 0);
                lerp = mrhe_blend_factors(lerp, m.fixed != 0);
        const float2* tab = reinterpret_cast<const float2*>(m.tables) + (size_t)level * m.T;
        float f0[8], f1[8];
#pragma unroll
        for (int c = 0; c < 8; c++) {
            float2 v = __ldg(tab + idx[c]);
            f0[c] = v.x;
            f1[c] = v.y;
        }
        out.x = mrhe_blend(f0, lerp);
        out.y = mrhe_blend(f1, lerp);
    }
    if (feat) reinterpret_cast<float2*>(feat)[q * 16 + level] = out;
    if (idxOut) {
        uint4* o = reinterpret_cast<uint4*>(idxOut + q * 128 + level * 8);
        o[0] = make_uint4(idx[0], idx[1], idx[2], idx[3]);
        o[1] = make_uint4(idx[4], idx[5], idx[6], idx[7]);
    }
}

__global__ void __launch_bounds__(256) k_encode_dir(const float* __restrict__ dir, size_t n, float* __restrict__ feat)
{
    size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    f3 d = mk3(__ldg(dir + 3 * q), __ldg(dir + 3 * q + 1), __ldg(dir + 3 * q + 2));
    float o[32];
    oneblob_encode(d, o, 1);
    float4* dst = reinterpret_cast<float4*>(feat + q * 32);
#pragma unroll
    for (int i = 0; i < 8; i++) dst[i] = make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
}

MrheDev make_mrhe_dev(const nrc_mrhe* m)
{
    MrheDev d;
    d.tables = m->tables;
    d.delta = m->delta;
    d.L = m->L;
    d.T = m->T;
    d.fixed = m->fixed ? 1u : 0u;
    for (int i = 0; i < 16; i++) d.res[i] = m->res[i];
    d.skySize[0] = 62.5f; d.skySize[1] = 42.5f; d.skySize[2] = 76.5f;   // vec3(125,85,153)/2, nrc-train.comp:212
    return d;
}

nrc_status launch_encode_pos(nrc_mrhe* m, const float* pos, size_t n, float* feat, uint32_t* idx)
{
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2, "the encoder kernels are written for featureCount == 2 (the reference's only value)");
    size_t threads = n * 16;
    unsigned blocks = (unsigned)((threads + 255) / 256);
    NRC_LAUNCH(m->ctx, m->ctx->stream, K_ENCODE_POS, k_encode_pos<<<blocks, 256, 0, m->ctx->stream>>>(pos, n, make_mrhe_dev(m), feat, idx));
    return NRC_OK;
}

nrc_status launch_encode_dir(nrc_ctx* ctx, const float* dir, size_t n, float* feat)
{
    if (n == 0) return NRC_OK;
    unsigned blocks = (unsigned)((n + 255) / 256);
    NRC_LAUNCH(ctx, ctx->stream, K_ENCODE_DIR, k_encode_dir<<<blocks, 256, 0, ctx->stream>>>(dir, n, feat));
    return NRC_OK;
}
