// scene.cu -- scene objects of the C-ABI: density volume, HDR environment map (+ its inverse CDFs).
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "trace.cuh"

// ---- volume (vk::Texture3D + en::VolumeData) -----------------------------------------------------------------------
extern "C" nrc_status nrc_volume_create(nrc_ctx* ctx, const float* density, uint32_t nx, uint32_t ny, uint32_t nz, nrc_volume** out)
{
    NRC_REQUIRE(ctx && density && out, "null argument");
    NRC_REQUIRE(nx > 0 && ny > 0 && nz > 0, "empty volume");
    size_t cells = (size_t)(nx + 1) * (ny + 1) * (nz + 1);
    NRC_REQUIRE(cells < ((size_t)1 << 31), "volume too large for 32-bit cell indices");
    NRC_CUDA(cudaSetDevice(ctx->device));
    nrc_volume* v = new nrc_volume();
    v->ctx = ctx; v->nx = nx; v->ny = ny; v->nz = nz;
    // defaults of VolumeData (src/VolumeData.cpp:82-89)
    memset(&v->params, 0, sizeof(v->params));
    v->params.densityFactor = 0.4f; v->params.g = 0.7f; v->params.noNnSpp = 1; v->params.withNnSpp = 1;
    NRC_CUDA(cudaMalloc(&v->bricks, cells * sizeof(uint2)));
    const void* dgrid;
    NRC_TRY(nrc_stage_in(ctx, 5, density, (size_t)nx * ny * nz * sizeof(float), &dgrid));
    NRC_TRY(launch_build_bricks(v, (const float*)dgrid));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = v;
    return NRC_OK;
}

extern "C" nrc_status nrc_volume_destroy(nrc_volume* v)
{
    if (!v) return NRC_OK;
    cudaStreamSynchronize(v->ctx->stream);
    cudaFree(v->bricks);
    delete v;
    return NRC_OK;
}

extern "C" nrc_status nrc_volume_set_params(nrc_volume* v, const nrc_volume_params* p)
{
    NRC_REQUIRE(v && p, "null argument");
    NRC_REQUIRE(p->noNnSpp >= 1 && p->withNnSpp >= 1 && p->noNnSpp <= 64 && p->withNnSpp <= 64, "spp must be 1..64");
    v->params = *p;
    return NRC_OK;
}

extern "C" nrc_status nrc_volume_sample(nrc_volume* v, const float* pos, size_t n, float* density)
{
    NRC_REQUIRE(v && pos && density, "null argument");
    if (n == 0) return NRC_OK;
    const void* dpos; void* dout;
    NRC_TRY(nrc_stage_in(v->ctx, 0, pos, n * 3 * sizeof(float), &dpos));
    NRC_TRY(nrc_stage_out(v->ctx, 3, density, n * sizeof(float), &dout));
    NRC_TRY(launch_volume_sample(v, (const float*)dpos, n, (float*)dout));
    NRC_TRY(nrc_stage_out_finish(v->ctx, density, dout, n * sizeof(float)));
    return NRC_OK;
}

// ---- environment map (en::HdrEnvMap) -------------------------------------------------------------------------------------
static nrc_status upload(nrc_ctx* ctx, void** dst, const void* src, size_t bytes)
{
    NRC_CUDA(cudaMalloc(dst, bytes));
    cudaMemcpyKind k = nrc_is_device_ptr(src) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    NRC_CUDA(cudaMemcpyAsync(*dst, src, bytes, k, ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    return NRC_OK;
}

extern "C" nrc_status nrc_envmap_create(nrc_ctx* ctx, uint32_t width, uint32_t height, const float* rgba, const float* invCdfX,
                                        const float* invCdfY, nrc_envmap** out)
{
    NRC_REQUIRE(ctx && rgba && out && width > 0 && height > 0, "bad argument");
    NRC_CUDA(cudaSetDevice(ctx->device));
    nrc_envmap* e = new nrc_envmap();
    e->ctx = ctx; e->w = width; e->h = height;
    NRC_TRY(upload(ctx, (void**)&e->rgba, rgba, (size_t)width * height * 4 * sizeof(float)));
    if (invCdfX) NRC_TRY(upload(ctx, (void**)&e->invCdfX, invCdfX, (size_t)width * height * sizeof(float)));
    if (invCdfY) NRC_TRY(upload(ctx, (void**)&e->invCdfY, invCdfY, (size_t)height * sizeof(float)));
    *out = e;
    return NRC_OK;
}

extern "C" nrc_status nrc_envmap_destroy(nrc_envmap* e)
{
    if (!e) return NRC_OK;
    cudaStreamSynchronize(e->ctx->stream);
    cudaFree(e->rgba);
    if (e->invCdfX) cudaFree(e->invCdfX);
    if (e->invCdfY) cudaFree(e->invCdfY);
    delete e;
    return NRC_OK;
}

extern "C" nrc_status nrc_envmap_set_strengths(nrc_envmap* e, float directStrength, float hpmStrength)
{
    NRC_REQUIRE(e, "null envmap");
    e->directStrength = directStrength; e->hpmStrength = hpmStrength;
    return NRC_OK;
}

// Opt-in estimator: `halfSampleCount = sampleCount / 2` as the author left it commented out at
// nrc-forward.frag:715 -- 4 of the 8 env directions per vertex come from the inverse CDFs of the map (fwd.frag:727-744).
// Consumes the same 16 random numbers per vertex, so the walks (and every other pixel's RNG stream) are unchanged.
extern "C" nrc_status nrc_envmap_set_importance_sampling(nrc_envmap* e, int on)
{
    NRC_REQUIRE(e, "null envmap");
    NRC_REQUIRE(!on || (e->invCdfX && e->invCdfY), "the env map was created without inverse CDFs (nrc_hdr4f_to_cdf)");
    e->cdfHalf = on != 0;
    return NRC_OK;
}

// ---- Hdr4fToCdf / InvertCdf (src/read_file.cpp:123-206) --------------------------------------------------------------------
// The reference sums each row sequentially in float; the running sums are kept sequential here too (one thread per
// row, strict adds) so the tables are bit-identical -- a parallel scan would round differently.  Rows are independent,
// which is all the parallelism a 512-row map needs.
__device__ void invert_cdf(const float* cdf, int n, float* inv)
{
    int p = 0;
    for (int i = 0; i < n; i++) {
        float threshold = __fdiv_rn((float)i, (float)n);
        while (p + 1 < n && cdf[p] < threshold) p++;
        inv[i] = __fdiv_rn((float)p, (float)n);
    }
}

__global__ void k_cdf_rows(const float4* __restrict__ rgba, int w, int h, float* __restrict__ cdfX, float* __restrict__ invCdfX,
                           float* __restrict__ pdfY)
{
    int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= h) return;
    float sum = 0.0f;
    float* row = cdfX + (size_t)y * w;
    for (int x = 0; x < w; x++) {
        float4 t = rgba[(size_t)y * w + x];
        float brightness = __fadd_rn(__fadd_rn(t.x, t.y), t.z);
        sum = __fadd_rn(sum, brightness);
        row[x] = sum;
    }
    for (int x = 0; x < w; x++) row[x] = __fdiv_rn(row[x], sum);
    pdfY[y] = sum;
    invert_cdf(row, w, invCdfX + (size_t)y * w);
}

__global__ void k_cdf_y(const float* __restrict__ pdfY, int h, float* __restrict__ cdfY, float* __restrict__ invCdfY)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    float sum = 0.0f;
    for (int y = 0; y < h; y++) { sum = __fadd_rn(sum, pdfY[y]); cdfY[y] = sum; }
    float last = cdfY[h - 1];
    for (int y = 0; y < h; y++) cdfY[y] = __fdiv_rn(cdfY[y], last);
    invert_cdf(cdfY, h, invCdfY);
}

extern "C" nrc_status nrc_hdr4f_to_cdf(nrc_ctx* ctx, const float* rgba, uint32_t width, uint32_t height, float* invCdfX, float* invCdfY)
{
    NRC_REQUIRE(ctx && rgba && invCdfX && invCdfY && width > 0 && height > 0, "bad argument");
    size_t px = (size_t)width * height;
    const void* drgba; void *dX, *dY, *tmp;
    NRC_TRY(nrc_stage_in(ctx, 0, rgba, px * 4 * sizeof(float), &drgba));
    NRC_TRY(nrc_stage_out(ctx, 3, invCdfX, px * sizeof(float), &dX));
    NRC_TRY(nrc_stage_out(ctx, 4, invCdfY, height * sizeof(float), &dY));
    NRC_TRY(nrc_scratch(ctx, 5, (px + 2 * (size_t)height) * sizeof(float), &tmp));
    float* cdfX = (float*)tmp;
    float* pdfY = cdfX + px;
    float* cdfY = pdfY + height;
    NRC_LAUNCH(ctx, ctx->stream, K_CDF_ROWS,
               k_cdf_rows<<<(height + 63) / 64, 64, 0, ctx->stream>>>((const float4*)drgba, (int)width, (int)height, cdfX, (float*)dX, pdfY));
    NRC_LAUNCH(ctx, ctx->stream, K_CDF_Y, k_cdf_y<<<1, 32, 0, ctx->stream>>>(pdfY, (int)height, cdfY, (float*)dY));
    NRC_TRY(nrc_stage_out_finish(ctx, invCdfX, dX, px * sizeof(float)));
    NRC_TRY(nrc_stage_out_finish(ctx, invCdfY, dY, height * sizeof(float)));
    return NRC_OK;
}
