// oracle_capi.cpp -- TEST INFRASTRUCTURE.  extern "C" surface of the CPU oracle for ctypes.
// PARITY UNPINNED (see nrc_oracle.hpp).  Nothing in the product may link or load this.
//
// Besides the four shader programs (nrc_oracle.hpp) this file restates the reference's few
// CPU-side helpers on the path:
//   src/Texture3D.cpp:25-40      float density[x][y][z] -> RGBA8 texels, x fastest
//   src/read_file.cpp:50-78      raw float32 grid file order  x*Y*Z + y*Z + z
//   src/read_file.cpp:123-206    InvertCdf / Hdr4fToCdf
//   src/MRHE.cpp:111-121         geometric resolution table
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "nrc_oracle.hpp"

using namespace glsl::oracle_impl;
using glsl::vec2; using glsl::vec3; using glsl::vec4; using glsl::uvec3; using glsl::mat4;

extern "C" {

void* orc_scene_create() { return new Scene(); }
void orc_scene_destroy(void* s) { delete (Scene*)s; }

void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

static mat4 mat_from16(const float* m)
{
    mat4 r;
    for (int c = 0; c < 4; c++) for (int rr = 0; rr < 4; rr++) r.c[c][rr] = m[c * 4 + rr];   // column-major like glm
    return r;
}

void orc_set_camera(void* sp, const float* projView, const float* invProjView, const float* pos)
{
    Scene& S = *(Scene*)sp;
    S.camMat.projView = mat_from16(projView);
    S.camMat.invProjView = mat_from16(invProjView);
    S.camera.pos = vec3(pos[0], pos[1], pos[2]);
}

// density is in the raw-file order of ReadFileDensity3D: index = x*ny*nz + y*nz + z.
void orc_set_volume(void* sp, const float* density, int nx, int ny, int nz)
{
    Scene& S = *(Scene*)sp;
    S.densityStorage.assign((size_t)nx * ny * nz * 4, 0);
    for (int i = 0; i < nx; i++)
        for (int j = 0; j < ny; j++)
            for (int k = 0; k < nz; k++) {
                float v = density[(size_t)i * ny * nz + (size_t)j * nz + k];
                uint8_t value = (uint8_t)(v * 255.0f);                                  // Texture3D.cpp:32
                size_t index = 4 * (size_t)i + 4 * (size_t)nx * j + 4 * (size_t)nx * ny * k;   // Texture3D.cpp:33
                S.densityStorage[index + 0] = value;
                S.densityStorage[index + 1] = value;
                S.densityStorage[index + 2] = value;
                S.densityStorage[index + 3] = 1;
            }
    S.densityTex.w = nx; S.densityTex.h = ny; S.densityTex.d = nz;
    S.densityTex.rgba = S.densityStorage.data();
}

void orc_set_volume_params(void* sp, const float* random4, unsigned useNN, unsigned showNonNN, float densityFactor,
                           float g, int noNnSpp, int withNnSpp)
{
    Scene& S = *(Scene*)sp;
    S.volumeData.random = vec4(random4[0], random4[1], random4[2], random4[3]);
    S.volumeData.useNN = useNN; S.volumeData.showNonNN = showNonNN;
    S.volumeData.densityFactor = densityFactor; S.volumeData.g = g;
    S.volumeData.noNnSpp = noNnSpp; S.volumeData.withNnSpp = withNnSpp;
}

void orc_set_dir_light(void* sp, const float* color, float zenith, const float* dir, float azimuth, float strength)
{
    Scene& S = *(Scene*)sp;
    S.dir_light.color = vec3(color[0], color[1], color[2]);
    S.dir_light.zenith = zenith; S.dir_light.azimuth = azimuth;
    S.dir_light.dir = vec3(dir[0], dir[1], dir[2]);
    S.dir_light.strength = strength;
}

void orc_set_point_light(void* sp, const float* pos, float strength, const float* color)
{
    Scene& S = *(Scene*)sp;
    S.pointLight.pos = vec3(pos[0], pos[1], pos[2]);
    S.pointLight.strength = strength;
    S.pointLight.color = vec3(color[0], color[1], color[2]);
}

void orc_set_envmap(void* sp, int w, int h, const float* rgba, const float* invCdfX, const float* invCdfY,
                    float directStrength, float hpmStrength)
{
    Scene& S = *(Scene*)sp;
    S.envStorage.assign(rgba, rgba + (size_t)w * h * 4);
    S.hdrEnvMap.w = w; S.hdrEnvMap.h = h; S.hdrEnvMap.ch = 4; S.hdrEnvMap.data = S.envStorage.data();
    if (invCdfX) {
        S.cdfXStorage.assign(invCdfX, invCdfX + (size_t)w * h);
        S.hdrEnvMapInvCdfX.w = w; S.hdrEnvMapInvCdfX.h = h; S.hdrEnvMapInvCdfX.ch = 1; S.hdrEnvMapInvCdfX.data = S.cdfXStorage.data();
    }
    if (invCdfY) {
        S.cdfYStorage.assign(invCdfY, invCdfY + (size_t)h);
        S.hdrEnvMapInvCdfY.w = h; S.hdrEnvMapInvCdfY.data = S.cdfYStorage.data();
    }
    S.hdrEnvMapData.directStrength = directStrength; S.hdrEnvMapData.hpmStrength = hpmStrength;
}

void orc_set_train_env_light(void* sp, int on) { ((Scene*)sp)->trainEnvLight = on != 0; }
void orc_set_env_cdf_half(void* sp, int on) { ((Scene*)sp)->envCdfHalf = on != 0; }
// the opt-in "fixed" preset
void orc_set_mrhe_fixed(void* sp, int on) { ((Scene*)sp)->mrheFixed = on != 0; }
void orc_set_adam(void* sp, int on, float beta2, float eps, float mrheBeta1)
{
    Scene* S = (Scene*)sp;
    S->adam = on != 0; S->adamBeta2 = beta2; S->adamEps = eps; S->mrheBeta1 = mrheBeta1;
    S->nrcSteps = S->mrheSteps = 0;
    S->secondMoment.clear(); S->mrFirstMoment.clear(); S->mrSecondMoment.clear();
}

void orc_set_mlp(void* sp, const float* params, float lr, float wd, float beta1)
{
    Scene& S = *(Scene*)sp;
    if (params) S.params.assign(params, params + kParamCount);
    S.nrcConfig.learningRate = lr; S.nrcConfig.weightDecay = wd; S.nrcConfig.beta1 = beta1;
}

void orc_set_mlp_state(void* sp, const float* deltaParams, const float* momentum)
{
    Scene& S = *(Scene*)sp;
    if (deltaParams) S.deltaParams.assign(deltaParams, deltaParams + kParamCount);
    if (momentum) S.momentum.assign(momentum, momentum + kParamCount);
}

void orc_get_mlp(void* sp, float* params, float* deltaParams, float* momentum)
{
    Scene& S = *(Scene*)sp;
    if (params) memcpy(params, S.params.data(), kParamCount * 4);
    if (deltaParams) memcpy(deltaParams, S.deltaParams.data(), kParamCount * 4);
    if (momentum) memcpy(momentum, S.momentum.data(), kParamCount * 4);
}

void orc_set_mrhe(void* sp, float lr, float wd, unsigned L, unsigned T, unsigned F, unsigned minRes, unsigned maxRes,
                  const unsigned* resolutions, const float* tables)
{
    Scene& S = *(Scene*)sp;
    S.mrhe.learningRate = lr; S.mrhe.weightDecay = wd; S.mrhe.levelCount = L; S.mrhe.hashTableSize = T;
    S.mrhe.featureCount = F; S.mrhe.minRes = minRes; S.mrhe.maxRes = maxRes;
    for (int i = 0; i < 16; i++) S.mrhe.resolutions[i] = resolutions[i];
    size_t n = (size_t)L * T * F;
    if (tables) S.mrHashTable.assign(tables, tables + n); else S.mrHashTable.assign(n, 0.0f);
    S.mrDeltaHashTable.assign(n, 0.0f);   // [A24] parity preset: zero deltas
}

void orc_set_mrhe_delta(void* sp, const float* delta)
{
    Scene& S = *(Scene*)sp;
    S.mrDeltaHashTable.assign(delta, delta + S.mrDeltaHashTable.size());
}

void orc_get_mrhe(void* sp, float* tables, float* delta)
{
    Scene& S = *(Scene*)sp;
    if (tables) memcpy(tables, S.mrHashTable.data(), S.mrHashTable.size() * 4);
    if (delta) memcpy(delta, S.mrDeltaHashTable.data(), S.mrDeltaHashTable.size() * 4);
}

void orc_set_train_size(void* sp, unsigned trainW, unsigned trainH)
{
    Scene& S = *(Scene*)sp;
    S.WIDTH_FACTOR = 1.0f / (float)trainW;    // NrcHpmRenderer.cpp:409-411
    S.HEIGHT_FACTOR = 1.0f / (float)trainH;
}

void orc_set_loss_scale(void* sp, float widthFactor, float heightFactor)
{
    Scene& S = *(Scene*)sp;
    S.WIDTH_FACTOR = widthFactor; S.HEIGHT_FACTOR = heightFactor;
}

float orc_get_mse_loss(void* sp) { return ((Scene*)sp)->mseLoss; }
void orc_reset_stats(void* sp) { ((Scene*)sp)->mseLoss = 0.0f; }

// MRHE.cpp:111-121 (float32 exp/log/pow of the C library, as the reference).
void orc_mrhe_resolutions(unsigned minRes, unsigned maxRes, unsigned levelCount, unsigned* out16)
{
    float b = std::exp((std::log((float)maxRes) - std::log((float)minRes)) / (float)(levelCount - 1));
    for (unsigned i = 0; i < 16; i++) out16[i] = 0;
    for (unsigned i = 0; i < levelCount && i < 16; i++) {
        float resF = (float)minRes * std::pow(b, (float)i);
        out16[i] = (unsigned)resF;
    }
}

// ---- stand-alone ops -------------------------------------------------------------------------

// EncodePosMrhe for N positions: feat[N][32], idx[N][128] (level*8 + neighbour), lerp[N][16][3].
void orc_encode_pos(void* sp, const float* pos, long N, float* feat, unsigned* idx, float* lerp)
{
    const Scene& S = *(Scene*)sp;
#pragma omp parallel for schedule(static)
    for (long n = 0; n < N; n++) {
        Invocation inv(S);
        inv.EncodePosMrhe(vec3(pos[3 * n], pos[3 * n + 1], pos[3 * n + 2]));
        if (feat) for (int i = 0; i < 32; i++) feat[32 * n + i] = inv.mrheFeatures[i];
        if (idx) for (int i = 0; i < 128; i++) idx[128 * n + i] = inv.allNeighbourIndices[i];
        if (lerp) for (int i = 0; i < 16; i++) {
            lerp[48 * n + 3 * i] = inv.allLerpFactors[i].x;
            lerp[48 * n + 3 * i + 1] = inv.allLerpFactors[i].y;
            lerp[48 * n + 3 * i + 2] = inv.allLerpFactors[i].z;
        }
    }
}

void orc_encode_dir(void* sp, const float* dir, long N, float* feat)
{
    const Scene& S = *(Scene*)sp;
#pragma omp parallel for schedule(static)
    for (long n = 0; n < N; n++) {
        Invocation inv(S);
        inv.EncodeDirOneBlob(vec3(dir[3 * n], dir[3 * n + 1], dir[3 * n + 2]));
        for (int i = 0; i < 32; i++) feat[32 * n + i] = inv.oneBlobFeatures[i];
    }
}

// Forward() (fwd.frag:503-534) for N queries -> rgb[N][3].
void orc_infer(void* sp, const float* pos, const float* dir, long N, float* rgb)
{
    const Scene& S = *(Scene*)sp;
#pragma omp parallel for schedule(static)
    for (long n = 0; n < N; n++) {
        Invocation inv(S);
        vec3 c = inv.Forward(vec3(pos[3 * n], pos[3 * n + 1], pos[3 * n + 2]), vec3(dir[3 * n], dir[3 * n + 1], dir[3 * n + 2]));
        rgb[3 * n] = c.x; rgb[3 * n + 1] = c.y; rgb[3 * n + 2] = c.z;
    }
}

// Backprop() (train.comp:985-1028) over N records, accumulated into the scene's delta buffers
// and mseLoss.  Samples are processed in fixed chunks whose partial sums are added in chunk
// order, so the result does not depend on the thread count.  pred (optional) gets the
// pre-loss network outputs; dropped[n] (optional) = 1 where the NaN guard skipped the sample.
void orc_train_accumulate(void* sp, const float* pos, const float* dir, const float* target, long N, float* pred,
                          unsigned char* dropped)
{
    Scene& S = *(Scene*)sp;
    int nThreads = orc_num_threads();
    // per-task sample count: 1024 for big batches, smaller when that would leave threads idle (sums are grouped per task, so the
    // grouping -- like the GPU's per-CTA partials -- only moves fp32 rounding)
    long chunk = N / ((long)nThreads * 4);
    if (chunk > 1024) chunk = 1024;
    if (chunk < 32) chunk = 32;
    const long nChunks = (N + chunk - 1) / chunk;
    const size_t tableN = S.mrDeltaHashTable.size();
    const size_t lvl0 = S.mrheFixed ? tableN : (size_t)S.mrhe.hashTableSize * 2;   // [A10]: only the first 2*T floats are ever touched (parity preset)
    long group = nThreads * 4;
    std::vector<std::vector<float>> dP(group, std::vector<float>(kParamCount));
    std::vector<std::vector<float>> dT(group, std::vector<float>(lvl0 <= tableN ? lvl0 : tableN));
    std::vector<double> dL(group);
    for (long base = 0; base < nChunks; base += group) {
        long cnt = (nChunks - base < group) ? (nChunks - base) : group;
#pragma omp parallel for schedule(dynamic, 1)
        for (long c = 0; c < cnt; c++) {
            std::fill(dP[c].begin(), dP[c].end(), 0.0f);
            std::fill(dT[c].begin(), dT[c].end(), 0.0f);
            GradSink sink{dP[c].data(), dT[c].data(), 0.0};
            long lo = (base + c) * chunk, hi = lo + chunk < N ? lo + chunk : N;
            for (long n = lo; n < hi; n++) {
                Invocation inv(S);
                float p3[3];
                bool ok = inv.Backprop(vec3(target[3 * n], target[3 * n + 1], target[3 * n + 2]),
                                       vec3(pos[3 * n], pos[3 * n + 1], pos[3 * n + 2]),
                                       vec3(dir[3 * n], dir[3 * n + 1], dir[3 * n + 2]), sink, p3);
                if (pred) { pred[3 * n] = ok ? p3[0] : 0.0f; pred[3 * n + 1] = ok ? p3[1] : 0.0f; pred[3 * n + 2] = ok ? p3[2] : 0.0f; }
                if (dropped) dropped[n] = ok ? 0 : 1;
            }
            dL[c] = sink.mseLoss;
        }
        for (long c = 0; c < cnt; c++) {
            for (size_t i = 0; i < kParamCount; i++) S.deltaParams[i] += dP[c][i];
            for (size_t i = 0; i < dT[c].size(); i++) S.mrDeltaHashTable[i] += dT[c][i];
            S.mseLoss += (float)dL[c];
        }
    }
}

void orc_nrc_step(void* sp) { NrcStep(*(Scene*)sp); }
void orc_mrhe_step(void* sp) { MrheStep(*(Scene*)sp); }

// ---- frame programs --------------------------------------------------------------------------

// nrc-forward.{vert,frag} over a W x H frame (rows y0..y1-1 only; pass 0,H for the full frame).
// rgba[W*H*4] float (vec4 outColor); counters[W*H*5] = {randDraws, litVertices, densityTaps,
// walkSteps, sdfSteps}; query[W*H*8] = {pos.xyz, dir.xyz, weight, valid}.  Any output may be NULL.
void orc_render(void* sp, unsigned W, unsigned H, unsigned y0, unsigned y1, float* rgba, unsigned* counters, float* query)
{
    const Scene& S = *(Scene*)sp;
    // pixels, not rows, are the parallel unit: a one-row call (the bench's evenly spread sample) still uses every thread
#pragma omp parallel for collapse(2) schedule(dynamic, 16)
    for (long py = (long)y0; py < (long)y1; py++)
        for (long pxl = 0; pxl < (long)W; pxl++) {
            const unsigned px = (unsigned)pxl;
            Invocation inv(S);
            vec4 c = inv.RenderPixel(px, (unsigned)py, W, H);
            size_t p = (size_t)py * W + px;
            if (rgba) { rgba[4 * p] = c.x; rgba[4 * p + 1] = c.y; rgba[4 * p + 2] = c.z; rgba[4 * p + 3] = c.w; }
            if (counters) {
                counters[5 * p] = inv.cnt.randDraws; counters[5 * p + 1] = inv.cnt.litVertices;
                counters[5 * p + 2] = inv.cnt.densityTaps; counters[5 * p + 3] = inv.cnt.walkSteps;
                counters[5 * p + 4] = inv.cnt.sdfSteps;
            }
            if (query) {
                const Query& q = inv.lastQuery;
                float* o = query + 8 * p;
                o[0] = q.pos.x; o[1] = q.pos.y; o[2] = q.pos.z; o[3] = q.dir.x; o[4] = q.dir.y; o[5] = q.dir.z;
                o[6] = q.weight; o[7] = q.valid ? 1.0f : 0.0f;
            }
        }
}

// nrc-train.comp main() for invocations i0..i1-1 (i = y*trainW + x) WITHOUT the Backprop:
// records[n][9] = {pos, dir, target}; counters[n][5] as above.
void orc_train_trace(void* sp, unsigned trainW, unsigned trainH, long i0, long i1, float* records, unsigned* counters)
{
    const Scene& S = *(Scene*)sp;
    (void)trainH;
#pragma omp parallel for schedule(dynamic, 4)
    for (long i = i0; i < i1; i++) {
        unsigned x = (unsigned)(i % trainW), y = (unsigned)(i / trainW);
        Invocation inv(S);
        vec3 p, d, t;
        inv.TrainInvocation(x, y, p, d, t);
        float* o = records + 9 * (i - i0);
        o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = d.x; o[4] = d.y; o[5] = d.z; o[6] = t.x; o[7] = t.y; o[8] = t.z;
        if (counters) {
            unsigned* c = counters + 5 * (i - i0);
            c[0] = inv.cnt.randDraws; c[1] = inv.cnt.litVertices; c[2] = inv.cnt.densityTaps; c[3] = inv.cnt.walkSteps;
            c[4] = inv.cnt.sdfSteps;
        }
    }
}

// One whole reference frame (NrcHpmRenderer.cpp:571-730): train -> nrc-step -> mrhe-step -> render.
void orc_frame(void* sp, unsigned W, unsigned H, unsigned trainW, unsigned trainH, float* rgba)
{
    Scene& S = *(Scene*)sp;
    orc_set_train_size(sp, trainW, trainH);
    long N = (long)trainW * trainH;
    std::vector<float> rec((size_t)N * 9), pos((size_t)N * 3), dir((size_t)N * 3), tgt((size_t)N * 3);
    orc_train_trace(sp, trainW, trainH, 0, N, rec.data(), nullptr);
    for (long n = 0; n < N; n++)
        for (int k = 0; k < 3; k++) { pos[3 * n + k] = rec[9 * n + k]; dir[3 * n + k] = rec[9 * n + 3 + k]; tgt[3 * n + k] = rec[9 * n + 6 + k]; }
    orc_train_accumulate(sp, pos.data(), dir.data(), tgt.data(), N, nullptr, nullptr);
    NrcStep(S);
    MrheStep(S);
    orc_render(sp, W, H, 0, H, rgba, nullptr, nullptr);
}

// ---- CPU-side helpers of the reference ---------------------------------------------------------

// read_file.cpp:123-139
static std::vector<float> InvertCdf(const std::vector<float>& cdf)
{
    std::vector<float> invCdf(cdf.size());
    size_t p = 0;
    for (size_t i = 0; i < invCdf.size(); i++) {
        const float threshold = (float)i / (float)invCdf.size();
        while (p + 1 < cdf.size() && cdf[p] < threshold) p++;   // (+ bound check: the reference reads past the end if rounding leaves cdf.back() < threshold)
        invCdf[i] = (float)p / (float)cdf.size();
    }
    return invCdf;
}

// read_file.cpp:141-206 -> invCdfX[h*w], invCdfY[h]
void orc_hdr4f_to_cdf(const float* hdr4f, long width, long height, float* invCdfXOut, float* invCdfYOut)
{
    std::vector<std::vector<float>> cdfX(height, std::vector<float>(width));
    std::vector<float> pdfY(height);
    for (long y = 0; y < height; y++) {
        float brightnessSum = 0.0f;
        for (long x = 0; x < width; x++) {
            const float r = hdr4f[(y * width * 4) + (x * 4) + 0];
            const float g = hdr4f[(y * width * 4) + (x * 4) + 1];
            const float b = hdr4f[(y * width * 4) + (x * 4) + 2];
            const float brightness = r + g + b;
            brightnessSum += brightness;
            cdfX[y][x] = brightnessSum;
        }
        for (long x = 0; x < width; x++) cdfX[y][x] /= brightnessSum;
        pdfY[y] = brightnessSum;
    }
    std::vector<float> cdfY(height);
    float brightnessSum = 0.0f;
    for (long y = 0; y < height; y++) { brightnessSum += pdfY[y]; cdfY[y] = brightnessSum; }
    const float last = cdfY[height - 1];
    for (long y = 0; y < height; y++) cdfY[y] /= last;   // (the reference divides by the running cdfY[height-1]; same value until the last element, which becomes 1 either way)
    for (long y = 0; y < height; y++) {
        std::vector<float> inv = InvertCdf(cdfX[y]);
        for (long x = 0; x < width; x++) invCdfXOut[(y * width) + x] = inv[x];
    }
    std::vector<float> invY = InvertCdf(cdfY);
    for (long y = 0; y < height; y++) invCdfYOut[y] = invY[y];
}

// getDensity (fwd.frag:579-582) at N world positions
void orc_density(void* sp, const float* pos, long N, float* out)
{
    const Scene& S = *(Scene*)sp;
    for (long n = 0; n < N; n++) {
        Invocation inv(S);
        out[n] = inv.getDensity(vec3(pos[3 * n], pos[3 * n + 1], pos[3 * n + 2]));
    }
}

// ---- nrc_math.h probes (tests/test_nrc_math.py) ----
void orc_math_eval(int fn, const float* x, const float* y, long N, float* out)
{
    for (long i = 0; i < N; i++) {
        switch (fn) {
            case 0: out[i] = nrc_sin(x[i]); break;
            case 1: out[i] = nrc_cos(x[i]); break;
            case 2: out[i] = nrc_exp(x[i]); break;
            case 3: out[i] = nrc_acos(x[i]); break;
            case 4: out[i] = nrc_asin(x[i]); break;
            case 5: out[i] = nrc_atan2(y[i], x[i]); break;
            case 6: out[i] = nrc_pow15(x[i]); break;
            case 7: { Scene S; Invocation inv(S); out[i] = inv.rand(vec2(x[i], y[i])); } break;
            default: out[i] = 0.0f;
        }
    }
}

unsigned orc_hash(unsigned x, unsigned y, unsigned z, unsigned T)
{
    Scene S; S.mrhe.hashTableSize = T;
    Invocation inv(S);
    return inv.HashFunc(uvec3(x, y, z));
}

}  // extern "C"
