// nrc_oracle.hpp -- TEST INFRASTRUCTURE.  CPU restatement of the reference's four GPU programs.
//
// PARITY UNPINNED: the reference (MichalSzczekocki/NeuralRadianceCaching) has no tests, no
// golden vectors and no CPU implementation of this path; its arithmetic
// is GLSL that needs a Vulkan driver, which neither this container nor the GPU box has.  This
// oracle is therefore a function-for-function restatement of the shader SOURCE, reviewed
// against it line by line; the known-answer values derivable from the source alone (hash
// values, resolution table, RNG fixed points, parameter counts) are pinned in
// tests/test_oracle_kat.py.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use anything under oracle/.
//
// Each function cites the shader lines it follows:
//   train.comp = data/shader/nrc-train/nrc-train.comp
//   fwd.frag   = data/shader/nrc-forward/nrc-forward.frag   (fwd.vert = nrc-forward.vert)
//   nrc-step   = data/shader/nrc-step/nrc-step.comp
//   mrhe-step  = data/shader/mrhe-step/mrhe-step.comp
// Quirks of the reference are reproduced on purpose and
// tagged [A<n>] where they occur.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>
#include "glsl_shim.hpp"

// Nested inside glsl so that unqualified sin/abs/floor/... resolve to the shim (scope lookup stops
// at namespace glsl) and never to the C library overloads in the global namespace.
namespace glsl {
namespace oracle_impl {

// ---------------------------------------------------------------------------------------------
// Resource interface: the seven descriptor sets as plain structs.
// ---------------------------------------------------------------------------------------------
struct CamMat { mat4 projView, invProjView; };                        // set 0.0 (train.comp:6-10)
struct CameraPos { vec3 pos; };                                       // set 0.1 (train.comp:12-15)
struct VolumeDataU {                                                  // set 1.1 (train.comp:19-28)
    vec4 random;
    uint useNN = 0, showNonNN = 0;
    float densityFactor = 0.4f, g = 0.7f;
    int noNnSpp = 1, withNnSpp = 1;
};
struct DirLightU { vec3 color; float zenith = 0; vec3 dir; float azimuth = 0; float strength = 0; };  // set 2.0
struct PointLightU { vec3 pos; float strength = 0; vec3 color; };     // set 4.0 (train.comp:165-170)
struct HdrEnvMapU { float directStrength = 1.0f, hpmStrength = 8.0f; };  // set 5.3
struct MrheU {                                                        // set 6.0 (train.comp:184-194)
    float learningRate = 0.01f, weightDecay = 0.0f;
    uint levelCount = 16, hashTableSize = 16384, featureCount = 2, minRes = 16, maxRes = 512;
    uint resolutions[16] = {0};   // explicit input [A5]
};
struct NrcConfig { float learningRate = 0.001f, weightDecay = 0.0f, beta1 = 0.5f; };  // set 3.36

// MLP parameter arena: W0..W4 (64x64, [row=out][col=in]), W5 (3x64), b0..b4 (64), b5 (3).
static const int kLayerIn[6] = {64, 64, 64, 64, 64, 64};
static const int kLayerOut[6] = {64, 64, 64, 64, 64, 3};
static const size_t kWeightOffset[6] = {0, 4096, 8192, 12288, 16384, 20480};
static const size_t kBiasOffset[6] = {20672, 20736, 20800, 20864, 20928, 20992};
static const size_t kParamCount = 20995;

struct Scene {
    CamMat camMat;
    CameraPos camera;
    sampler3D densityTex;
    std::vector<uint8_t> densityStorage;   // RGBA8, as src/Texture3D.cpp:25-40 builds it
    VolumeDataU volumeData;
    DirLightU dir_light;
    PointLightU pointLight;
    sampler2D hdrEnvMap;
    sampler2D hdrEnvMapInvCdfX;
    sampler1D hdrEnvMapInvCdfY;
    std::vector<float> envStorage, cdfXStorage, cdfYStorage;
    HdrEnvMapU hdrEnvMapData;
    bool trainEnvLight = false;  // opt-in: the env-map term of TraceScene that train.comp:1118 has commented out
    bool envCdfHalf = false;   // opt-in: halfSampleCount = sampleCount / 2, the variant commented out at fwd.frag:715
    // opt-in "fixed" preset (none of it is what the reference ships):
    //   mrheFixed: grid-space lerp factors (resPos - floor(resPos)) instead of world-minus-grid [A2], each factor blends its own axis
    //              [A3], every feature gets its own error [A9], gradients land in their level's table [A10];
    //   adam:      second moment + bias correction in both step shaders (the shipped "Adam" is first-moment momentum only [A16]).
    bool mrheFixed = false;
    bool adam = false;
    float adamBeta2 = 0.999f, adamEps = 1e-8f, mrheBeta1 = 0.9f;
    uint32_t nrcSteps = 0, mrheSteps = 0;
    std::vector<float> secondMoment, mrFirstMoment, mrSecondMoment;
    MrheU mrhe;
    std::vector<float> mrHashTable, mrDeltaHashTable;
    NrcConfig nrcConfig;
    std::vector<float> params, deltaParams, momentum;   // kParamCount each
    float mseLoss = 0.0f;                                 // set 3.37
    float WIDTH_FACTOR = 0.01f, HEIGHT_FACTOR = 0.01f;    // spec constants (NrcHpmRenderer.cpp:409-411)
    vec3 skySize = vec3(125.0f, 85.0f, 153.0f) / 2.0f;    // train.comp:212
    vec3 skyPos = vec3(0.0f);                             // train.comp:213

    Scene() : params(kParamCount, 0.0f), deltaParams(kParamCount, 0.0f), momentum(kParamCount, 0.0f) {}
    const float* W(int l) const { return params.data() + kWeightOffset[l]; }
    const float* B(int l) const { return params.data() + kBiasOffset[l]; }
};

// Gradient sink: Backprop() adds here instead of doing global float atomics; the caller sums
// sinks in a fixed order (the reference's atomic order is itself unordered).
struct GradSink {
    float* deltaParams;        // kParamCount
    float* mrDeltaHashTable;   // L*T*F
    double mseLoss;            // accumulated like nrcStats.mseLoss
};

// Per-invocation diagnostic counters (not in the shaders; they are what "bit-exact sample
// counts" is checked on).
struct Counters {
    uint32_t randDraws = 0;      // number of myRand() calls
    uint32_t litVertices = 0;    // walk vertices with density > 0 that evaluated TraceScene
    uint32_t densityTaps = 0;    // getDensity() calls
    uint32_t walkSteps = 0;      // iterations of the walk loop(s) executed
    uint32_t sdfSteps = 0;       // sky_sdf evaluations inside find_entry_exit
};

struct Query { vec3 pos, dir; float weight; bool valid; };   // the NRC query a pixel terminated into

#define PI 3.14159265359f
#define MAX_RAY_DISTANCE 100000.0f
#define MIN_RAY_DISTANCE 0.125f

// One shader invocation: the GLSL file-scope mutable globals become members.
struct Invocation {
    const Scene& S;
    Counters cnt;
    float preRand = 0.0f, prePreRand = 0.0f;         // train.comp:223-224
    float mrheFeatures[32];                          // train.comp:262
    uint allNeighbourIndices[128];                   // train.comp:263
    vec3 allLerpFactors[16];                         // train.comp:264
    float oneBlobFeatures[32];                       // train.comp:342
    float nr0[64], nr1[64], nr2[64], nr3[64], nr4[64], nr5[64], nr6[4];   // train.comp:475-481 (+1 for [A7])
    Query lastQuery;

    explicit Invocation(const Scene& s) : S(s) { lastQuery.valid = false; lastQuery.weight = 0.0f; }

    // ---- Random (train.comp:226-243, fwd.frag:152-169) ----
    float rand(vec2 co) { return fract(sin(dot(co, vec2(12.9898f, 78.233f))) * 43758.5453f); }
    float myRand()
    {
        cnt.randDraws++;
        float result = rand(vec2(preRand, prePreRand));
        prePreRand = preRand;
        preRand = result;
        return result;
    }
    float RandFloat(float maxVal) { float f = myRand(); return f * maxVal; }

    // ---- MRHE (train.comp:247-339, fwd.frag:173-252) ----
    float GetMrheFeature(uint level, uint entryIndex, uint featureIndex) const
    {
        const uint linearIndex = (S.mrhe.hashTableSize * S.mrhe.featureCount * level) + (entryIndex * S.mrhe.featureCount) + featureIndex;
        return S.mrHashTable[linearIndex];
    }
    uint HashFunc(const uvec3 pos) const
    {
        const uvec3 primes = uvec3(1u, 19349663u, 83492791u);
        uint hash = (pos.x * primes.x) + (pos.y * primes.y) + (pos.z * primes.z);   // [A1] '+', wrapping
        hash %= S.mrhe.hashTableSize;
        return hash;
    }
    void EncodePosMrhe(const vec3 pos)
    {
        const vec3 normPos = (pos / S.skySize) + vec3(0.5f);
        for (uint level = 0; level < S.mrhe.levelCount; level++) {
            const uint res = S.mrhe.resolutions[level];
            const vec3 resPos = normPos * float(res);
            const vec3 floorPos = floor(resPos);
            vec3 neighbours[8];
            for (uint x = 0; x < 2; x++)
                for (uint y = 0; y < 2; y++)
                    for (uint z = 0; z < 2; z++) {
                        uint linearIndex = (x * 4) + (y * 2) + z;
                        neighbours[linearIndex] = floorPos + to_vec3(uvec3(x, y, z));
                    }
            uint neighbourIndices[8];
            for (uint neigh = 0; neigh < 8; neigh++) {
                const uint index = HashFunc(uvec3(neighbours[neigh]));   // [A4] saturating cast
                neighbourIndices[neigh] = index;
                allNeighbourIndices[(level * 8) + neigh] = index;
            }
            vec2 neighbourFeatures[8];
            for (uint neigh = 0; neigh < 8; neigh++) {
                const uint entryIndex = neighbourIndices[neigh];
                neighbourFeatures[neigh] = vec2(GetMrheFeature(level, entryIndex, 0), GetMrheFeature(level, entryIndex, 1));
            }
            vec3 lerpFactors = pos - neighbours[0];   // [A2] world pos minus grid-space floor
            if (S.mrheFixed) lerpFactors = resPos - floorPos;
            allLerpFactors[level] = lerpFactors;
            if (S.mrheFixed) lerpFactors = vec3(lerpFactors.z, lerpFactors.y, lerpFactors.x);   // the blend below pairs .z with the x bit [A3]
            vec2 zLerpFeatures[4];
            for (uint i = 0; i < 4; i++)              // [A3] "z" lerp blends the x bit
                zLerpFeatures[i] = (neighbourFeatures[i] * (1.0f - lerpFactors.z)) + (neighbourFeatures[4 + i] * lerpFactors.z);
            vec2 yLerpFeatures[2];
            for (uint i = 0; i < 2; i++)
                yLerpFeatures[i] = (zLerpFeatures[i] * (1.0f - lerpFactors.y)) + (zLerpFeatures[2 + i] * lerpFactors.y);
            vec2 xLerpFeatures = (yLerpFeatures[0] * (1.0f - lerpFactors.x)) + (yLerpFeatures[1] * lerpFactors.x);
            mrheFeatures[(level * S.mrhe.featureCount) + 0] = xLerpFeatures.x;
            mrheFeatures[(level * S.mrhe.featureCount) + 1] = xLerpFeatures.y;
        }
    }

    // ---- OneBlob (train.comp:344-365) ----
    static float NormGauss(const float x, const float m, const float sigma)
    {
        const float term1 = 1.0f / (sigma * sqrt(2.0f * PI));
        const float term2 = ((x - m) / sigma);
        return term1 * exp(-0.5f * term2 * term2);
    }
    void EncodeDirOneBlob(const vec3 dir)
    {
        const float theta = (atan(dir.z, dir.x) / PI) + 0.5f;
        const float phi = (atan(length(dir.xz()), dir.y) / PI) + 0.5f;
        const float sigma = 1.0f / 4.0f;
        for (uint i = 0; i < 16; i++) {
            const float fI = float(i);
            oneBlobFeatures[i] = NormGauss(fI, theta, sigma);
            oneBlobFeatures[i + 16] = NormGauss(fI, phi, sigma);
        }
    }

    // ---- path-trace helpers (train.comp:368-471, fwd.frag:537-639) ----
    float sky_sdf(vec3 pos)
    {
        cnt.sdfSteps++;
        vec3 d = abs(pos - S.skyPos) - S.skySize / 2.0f;
        return length(max(d, 0.0f)) + min(max(d.x, max(d.y, d.z)), 0.0f);
    }
    void find_entry_exit(vec3 ro, vec3 rd, vec3 out[2])
    {
        float dist;
        do {
            dist = sky_sdf(ro);
            ro += dist * rd;
        } while (dist > MIN_RAY_DISTANCE && dist < MAX_RAY_DISTANCE);
        vec3 entry = ro;
        ro += rd * length(2.0f * S.skySize);
        rd *= -1.0f;
        do {
            dist = sky_sdf(ro);
            ro += dist * rd;
        } while (dist > MIN_RAY_DISTANCE && dist < MAX_RAY_DISTANCE);
        out[0] = entry;
        out[1] = ro;
    }
    vec3 get_sky_uvw(vec3 pos) const { return ((pos - S.skyPos) / S.skySize) + vec3(0.5f); }
    float getDensity(vec3 pos)
    {
        cnt.densityTaps++;
        return S.volumeData.densityFactor * texture_r(S.densityTex, get_sky_uvw(pos));
    }
    float hg_phase_func(const float cos_theta) const
    {
        const float g = S.volumeData.g;
        const float g2 = g * g;
        return 0.5f * (1.0f - g2) / nrc_pow15(1.0f + g2 - (2.0f * g * cos_theta));
    }
    static mat4 rotationMatrix(vec3 axis, float angle)
    {
        axis = normalize(axis);
        float s = sin(angle);
        float c = cos(angle);
        float oc = 1.0f - c;
        return mat4(oc * axis.x * axis.x + c,          oc * axis.x * axis.y - axis.z * s, oc * axis.z * axis.x + axis.y * s, 0.0f,
                    oc * axis.x * axis.y + axis.z * s, oc * axis.y * axis.y + c,          oc * axis.y * axis.z - axis.x * s, 0.0f,
                    oc * axis.z * axis.x - axis.y * s, oc * axis.y * axis.z + axis.x * s, oc * axis.z * axis.z + c,          0.0f,
                    0.0f, 0.0f, 0.0f, 1.0f);
    }
    vec3 NewRayDir(vec3 oldRayDir)
    {
        oldRayDir = normalize(oldRayDir);
        vec3 orthoDir = oldRayDir.z < oldRayDir.x ? vec3(oldRayDir.y, -oldRayDir.x, 0.0f) : vec3(0.0f, -oldRayDir.z, oldRayDir.y);
        orthoDir = normalize(orthoDir);
        float g = S.volumeData.g;
        float cosTheta;
        if (abs(g) < 0.001f) {
            cosTheta = 1.0f - 2.0f * RandFloat(1.0f);
        } else {
            float sqrTerm = (1.0f - g * g) / (1.0f - g + (2.0f * g * RandFloat(1.0f)));
            cosTheta = (1.0f + (g * g) - (sqrTerm * sqrTerm)) / (2.0f * g);
        }
        float angle = acos(cosTheta);
        mat4 rotMat = rotationMatrix(orthoDir, angle);
        vec3 newRayDir = (rotMat * vec4(oldRayDir, 1.0f)).xyz();
        angle = RandFloat(2.0f * PI);
        rotMat = rotationMatrix(oldRayDir, angle);
        newRayDir = (rotMat * vec4(newRayDir, 1.0f)).xyz();
        return normalize(newRayDir);
    }

    // ---- NN forward (train.comp:547-741, fwd.frag:299-534) ----
    void ApplyWeights(int l, const float* in, float* out) const
    {
        const float* W = S.W(l);
        const float* b = S.B(l);
        for (int outRow = 0; outRow < kLayerOut[l]; outRow++) {
            float sum = 0.0f;
            for (int inCol = 0; inCol < 64; inCol++) sum += in[inCol] * W[outRow * 64 + inCol];
            out[outRow] = sum + b[outRow];
        }
    }
    static void Activate(float* v, int n) { for (int i = 0; i < n; i++) v[i] = max(0.0f, v[i]); }
    void EncodeRay(vec3 pos, const vec3 dir)
    {
        EncodePosMrhe(pos);
        EncodeDirOneBlob(dir);
        for (uint i = 0; i < 32; i++) { nr0[i] = mrheFeatures[i]; nr0[i + 32] = oneBlobFeatures[i]; }
    }
    void ForwardLayers()
    {
        ApplyWeights(0, nr0, nr1); Activate(nr1, 64);
        ApplyWeights(1, nr1, nr2); Activate(nr2, 64);
        ApplyWeights(2, nr2, nr3); Activate(nr3, 64);
        ApplyWeights(3, nr3, nr4); Activate(nr4, 64);
        ApplyWeights(4, nr4, nr5); Activate(nr5, 64);
        nr6[3] = 0.0f;
        ApplyWeights(5, nr5, nr6); Activate(nr6, 3);   // [A7] the 4th element is ignored
    }
    // fwd.frag:503-534
    vec3 Forward(vec3 ro, const vec3 rd)
    {
        EncodeRay(ro, rd);
        ForwardLayers();
        return vec3(max(0.0f, nr6[0]), max(0.0f, nr6[1]), max(0.0f, nr6[2]));
    }

    // ---- NN backward (train.comp:743-1028) ----
    static float ReluDeriv(float x) { return x > 0.0f ? 1.0f : 0.0f; }
    // One BackpropL(): err = layer-l output error (in place), lower = nr_l (the layer's input
    // activations), which is overwritten with the propagated error.  [A6] the gate uses lower[i].
    void BackpropLayer(int l, float* err, float* lower, GradSink& sink) const
    {
        const int nOut = kLayerOut[l];
        const float opc = S.WIDTH_FACTOR * S.HEIGHT_FACTOR;   // ONE_OVER_PIXEL_COUNT [A8]
        for (int i = 0; i < nOut; i++) err[i] *= ReluDeriv(lower[i]);
        float* dW = sink.deltaParams + kWeightOffset[l];
        float* dB = sink.deltaParams + kBiasOffset[l];
        for (int row = 0; row < nOut; row++) {
            for (int col = 0; col < 64; col++) {
                float deltaWeight = -lower[col] * err[row];
                dW[row * 64 + col] += deltaWeight * opc;
            }
            dB[row] += -err[row] * opc;
        }
        const float* W = S.W(l);
        float tmp[64];
        for (int col = 0; col < 64; col++) {
            float error = 0.0f;
            for (int row = 0; row < nOut; row++) error += W[row * 64 + col] * err[row];
            tmp[col] = error;
        }
        for (int col = 0; col < 64; col++) lower[col] = tmp[col];
    }
    void BackpropMrhe(GradSink& sink) const
    {
        const float opc = S.WIDTH_FACTOR * S.HEIGHT_FACTOR;
        for (uint level = 0; level < 16; level++) {
            const vec3 lerpFactors = allLerpFactors[level];
            const vec2 error = vec2(nr0[(level * 2) + 0], nr0[(level * 2) + (S.mrheFixed ? 1 : 0)]);   // [A9]
            const size_t levelOffset = S.mrheFixed ? (size_t)level * S.mrhe.hashTableSize * 2 : 0;   // [A10]
            for (uint x = 0; x < 2; x++)
                for (uint y = 0; y < 2; y++)
                    for (uint z = 0; z < 2; z++) {
                        const uint linearIndex = (x * 4) + (y * 2) + z;
                        const uint tableEntryIndex = allNeighbourIndices[(level * 8) + linearIndex];
                        const float xFactor = x == 1 ? lerpFactors.x : (1.0f - lerpFactors.x);   // [A11]
                        const float yFactor = y == 1 ? lerpFactors.y : (1.0f - lerpFactors.y);
                        const float zFactor = z == 1 ? lerpFactors.z : (1.0f - lerpFactors.z);
                        const float errorWeight = xFactor * yFactor * zFactor;
                        const vec2 delta = -error * errorWeight * opc;
                        sink.mrDeltaHashTable[levelOffset + (2 * tableEntryIndex) + 0] += delta.x;   // [A10] no level offset
                        sink.mrDeltaHashTable[levelOffset + (2 * tableEntryIndex) + 1] += delta.y;
                    }
        }
    }
    // train.comp:985-1028.  Returns false when the NaN guard dropped the sample [A12].
    bool Backprop(vec3 target, const vec3 pos, const vec3 dir, GradSink& sink, float pred_out[3] = nullptr)
    {
        target = min(target, vec3(1024.0f));
        EncodeRay(pos, dir);
        for (uint i = 0; i < 5; i++) if (isnan(nr0[i])) return false;
        ForwardLayers();
        const vec3 pred = vec3(nr6[0], nr6[1], nr6[2]);
        if (pred_out) { pred_out[0] = pred.x; pred_out[1] = pred.y; pred_out[2] = pred.z; }
        const vec3 error = pred - target;
        const float mse = ((error.x * error.x) + (error.y * error.y) + (error.z * error.z)) / 3.0f;
        sink.mseLoss += (double)(mse * (S.WIDTH_FACTOR * S.HEIGHT_FACTOR));
        nr6[0] = 2.0f * error.x; nr6[1] = 2.0f * error.y; nr6[2] = 2.0f * error.z;
        BackpropLayer(5, nr6, nr5, sink);
        BackpropLayer(4, nr5, nr4, sink);
        BackpropLayer(3, nr4, nr3, sink);
        BackpropLayer(2, nr3, nr2, sink);
        BackpropLayer(1, nr2, nr1, sink);
        BackpropLayer(0, nr1, nr0, sink);
        BackpropMrhe(sink);
        return true;
    }

    // ---- transmittance and lights (train.comp:1032-1120, fwd.frag:641-755) ----
    float GetTransmittance(const vec3 start, const vec3 end, const uint count)
    {
        const vec3 dir = end - start;
        const float stepSize = length(dir) / float(count);
        if (stepSize == 0.0f) return 1.0f;
        float transmittance = 1.0f;
        for (uint i = 0; i < count; i++) {
            const float factor = float(i) / float(count);
            const vec3 samplePoint = start + (factor * dir);
            const float density = getDensity(samplePoint);
            const float t_r = exp(-density * stepSize);
            transmittance *= t_r;
        }
        return transmittance;
    }
    vec3 TraceDirLight(const vec3 pos, const vec3 dir)
    {
        if (S.dir_light.strength == 0.0f) return vec3(0.0f);
        vec3 ee[2];
        find_entry_exit(pos, -normalize(S.dir_light.dir), ee);
        const float transmittance = GetTransmittance(pos, ee[1], 32);
        const float phase = hg_phase_func(dot(S.dir_light.dir, -dir));
        return vec3(1.0f) * transmittance * S.dir_light.strength * phase;   // [A21] colour ignored
    }
    vec3 TracePointLight(const vec3 pos, const vec3 dir)
    {
        if (S.pointLight.strength == 0.0f) return vec3(0.0f);
        const float transmittance = GetTransmittance(S.pointLight.pos, pos, 32);
        const float phase = hg_phase_func(dot(normalize(S.pointLight.pos - pos), -dir));
        return S.pointLight.color * S.pointLight.strength * transmittance * phase;
    }
    // fwd.frag:690-708
    vec3 SampleHdrEnvMapUV(const vec2 dir, const bool hpm) const
    {
        const vec2 invAtan = vec2(0.1591f, 0.3183f);   // [A22]
        vec2 uv = dir;
        uv = uv * invAtan;
        uv = uv + 0.5f;
        const float strength = hpm ? S.hdrEnvMapData.hpmStrength : S.hdrEnvMapData.directStrength;
        return texture(S.hdrEnvMap, uv).xyz() * strength;
    }
    vec3 SampleHdrEnvMapDir(const vec3 dir, const bool hpm) const
    {
        vec2 phiTheta = vec2(atan(dir.z, dir.x), asin(dir.y));
        return SampleHdrEnvMapUV(phiTheta, hpm);
    }
    // fwd.frag:710-749
    vec3 SampleHdrEnvMapGather(const vec3 pos, const vec3 dir, uint sampleCount)
    {
        vec3 light = vec3(0.0f);
        const uint halfSampleCount = S.envCdfHalf ? sampleCount / 2 : sampleCount;   // as shipped the CDF-importance half never runs
        for (uint i = 0; i < halfSampleCount; i++) {
            const vec3 randomDir = NewRayDir(dir);
            const float phase = 1.0f;
            vec3 ee[2];
            find_entry_exit(pos, randomDir, ee);
            const float transmittance = GetTransmittance(pos, ee[1], 16);
            const vec3 sampleLight = SampleHdrEnvMapDir(randomDir, true) * phase * transmittance;
            light += sampleLight;
        }
        for (uint i = 0; i < sampleCount - halfSampleCount; i++) {
            const float thetaNorm = texture(S.hdrEnvMapInvCdfY, RandFloat(1.0f)).x;
            const float phiNorm = texture(S.hdrEnvMapInvCdfX, vec2(RandFloat(1.0f), thetaNorm)).x;
            // As written (fwd.frag:735) the direction is sin(theta) * (cos(phi), 1, sin(phi)): not a unit vector (|d| up to 1.41 makes
            // hg_phase_func take pow of a negative number -> NaN) and not the direction of texel (phiNorm, thetaNorm) under the map's
            // own uv convention (uv = (atan(z,x), asin(y)) * (1/2pi, 1/pi) + 0.5, fwd.frag:690-708).  The loop is dead code in the
            // reference; the opt-in uses the unit direction that convention assigns to the texel.
            const float th = thetaNorm * PI, ph = phiNorm * 2.0f * PI;
            const vec3 randomDir = vec3(-sin(th) * cos(ph), -cos(th), -sin(th) * sin(ph));
            const float phase = hg_phase_func(dot(randomDir, -dir));
            vec3 ee[2];
            find_entry_exit(pos, randomDir, ee);
            const float transmittance = GetTransmittance(pos, ee[1], 16);
            const vec3 sampleLight = texture(S.hdrEnvMap, vec2(phiNorm, thetaNorm)).xyz() * S.hdrEnvMapData.hpmStrength * phase * transmittance;
            light += sampleLight;
        }
        light /= float(sampleCount);
        return light;
    }
    vec3 TraceSceneRender(const vec3 pos, const vec3 dir)   // fwd.frag:751-755
    {
        vec3 a = TraceDirLight(pos, dir);
        vec3 b = TracePointLight(pos, dir);
        vec3 c = SampleHdrEnvMapGather(pos, dir, 8);
        return a + b + c;
    }
    vec3 TraceSceneTrain(const vec3 pos, const vec3 dir)    // train.comp:1116-1120 (no env map)
    {
        vec3 a = TraceDirLight(pos, dir);
        vec3 b = TracePointLight(pos, dir);
        if (S.trainEnvLight) {
            vec3 c = SampleHdrEnvMapGather(pos, dir, 8);
            return a + b + c;
        }
        return a + b;
    }

    // ---- render path (fwd.frag:757-834) ----
    vec4 TracePathRender(const vec3 rayOrigin, const vec3 rayDir, bool useNN)
    {
        vec3 scatteredLight = vec3(0.0f);
        float transmittance = 1.0f;
        vec3 ee[2];
        find_entry_exit(rayOrigin, rayDir, ee);
        const vec3 entry = ee[0];
        vec3 currentPoint = entry;
        vec3 lastPoint = entry;
        vec3 currentDir = rayDir;
        vec3 lastDir = vec3(0.0f);
        float totalTermProb = 1.0f;
        for (uint i = 0; i < 32; i++) {
            cnt.walkSteps++;
            const float density = getDensity(currentPoint);
            if (density > 0.0f) {
                if (useNN) {
                    if (RandFloat(1.0f) > totalTermProb) {
                        if (S.volumeData.showNonNN == 0) {
                            const float dirPhase = hg_phase_func(dot(currentDir, -lastDir));   // [A19]
                            lastQuery.pos = currentPoint; lastQuery.dir = currentDir;
                            lastQuery.weight = transmittance * dirPhase; lastQuery.valid = true;
                            scatteredLight += transmittance * Forward(currentPoint, currentDir) * dirPhase;
                        }
                        return vec4(scatteredLight, transmittance);
                    }
                    totalTermProb *= 0.5f;
                }
                cnt.litVertices++;
                const vec3 sceneLighting = TraceSceneRender(currentPoint, currentDir);
                const float phase = 1.0f;
                const vec3 s_int = density * sceneLighting * phase;
                const float t_r = GetTransmittance(currentPoint, lastPoint, 32);
                scatteredLight += transmittance * s_int;
                transmittance *= t_r;
                lastPoint = currentPoint;
                lastDir = currentDir;
                currentDir = NewRayDir(currentDir);
            }
            find_entry_exit(currentPoint, currentDir, ee);
            const float maxDistance = distance(ee[1], currentPoint) * 0.1f;
            const float nextDistance = RandFloat(maxDistance);
            currentPoint = currentPoint + (currentDir * nextDistance);
        }
        return vec4(scatteredLight, transmittance);
    }
    vec4 TracePathMultiple(const vec3 rayOrigin, const vec3 rayDir, bool useNN)
    {
        const uint spp = (uint)(useNN ? S.volumeData.withNnSpp : S.volumeData.noNnSpp);
        vec4 average = vec4(0.0f);
        for (uint i = 0; i < spp; i++) average += TracePathRender(rayOrigin, rayDir, useNN);
        average /= float(spp);
        return average;
    }
    // fwd.vert:28-40 + fwd.frag:837-870 for the pixel (px, py) of a W x H frame; fragUV is the
    // pixel centre, pixelWorldPos the un-projection of the (y-flipped [A18]) screen coordinate.
    vec4 RenderPixel(uint px, uint py, uint W, uint H)
    {
        const vec2 fragUV = vec2((float(px) + 0.5f) / float(W), (float(py) + 0.5f) / float(H));
        vec4 screenCoord = vec4((fragUV.x * 2.0f) - 1.0f, (fragUV.y * 2.0f) - 1.0f, 0.0f, 1.0f);
        screenCoord.y *= -1.0f;
        const vec4 worldPos = S.camMat.invProjView * screenCoord;
        const vec3 pixelWorldPos = worldPos.xyz() / worldPos.w;
        preRand = S.volumeData.random.x * fragUV.x;       // fwd.frag:149-150
        prePreRand = S.volumeData.random.y * fragUV.y;
        const vec3 ro = S.camera.pos;
        const vec3 rd = normalize(pixelWorldPos - ro);
        vec3 ee[2];
        find_entry_exit(ro, rd, ee);
        vec3 envMapColor = SampleHdrEnvMapDir(rd, false);
        if (sky_sdf(ee[0]) > MAX_RAY_DISTANCE) return vec4(envMapColor, 1.0f);
        const vec4 traceResult = TracePathMultiple(ro, rd, S.volumeData.useNN == 1);
        if (traceResult.w == 1.0f) return vec4(envMapColor, 1.0f);   // [A20]
        return traceResult;
    }

    // ---- train path (train.comp:1122-1259) ----
    vec3 TracePathTrain(const vec3 rayOrigin, const vec3 rayDir)
    {
        vec3 scatteredLight = vec3(0.0f);
        float transmittance = 1.0f;
        vec3 ee[2];
        find_entry_exit(rayOrigin, rayDir, ee);
        const vec3 entry = ee[0];
        vec3 currentPoint = entry;
        vec3 lastPoint = entry;
        vec3 currentDir = rayDir;
        for (uint i = 0; i < 64; i++) {
            cnt.walkSteps++;
            const float density = getDensity(currentPoint);
            if (density > 0.0f) {
                cnt.litVertices++;
                const vec3 sceneLighting = TraceSceneTrain(currentPoint, currentDir);
                const vec3 s_int = density * sceneLighting;
                const float t_r = GetTransmittance(currentPoint, lastPoint, 32);
                scatteredLight += transmittance * s_int;
                transmittance *= t_r;
                lastPoint = currentPoint;
                currentDir = NewRayDir(currentDir);
            }
            find_entry_exit(currentPoint, currentDir, ee);
            const float maxDistance = distance(ee[1], currentPoint) * 0.1f;
            const float nextDistance = RandFloat(maxDistance);
            currentPoint = currentPoint + (currentDir * nextDistance);
        }
        return scatteredLight;
    }
    // train.comp:1173-1226 without the final Backprop: returns the training record.
    void TracePathAndRecord(const vec3 rayOrigin, const vec3 rayDir, vec3& outPos, vec3& outDir, vec3& outTarget)
    {
        vec3 ee[2];
        find_entry_exit(rayOrigin, rayDir, ee);
        const vec3 entry = ee[0];
        vec3 currentPoint = entry;
        vec3 currentDir = rayDir;
        float totalTermProb = 1.0f;
        bool didScatter = false;
        for (uint i = 0; i < 64; i++) {
            if (RandFloat(1.0f) > totalTermProb) break;
            totalTermProb *= 0.5f;
            cnt.walkSteps++;
            if (getDensity(currentPoint) > 0.0f) {
                didScatter = true;
                currentDir = NewRayDir(currentDir);
            }
            vec3 ee2[2];
            find_entry_exit(currentPoint, currentDir, ee2);
            const float maxDistance = distance(ee2[1], currentPoint) * 0.1f;
            const float nextDistance = maxDistance * RandFloat(1.0f);
            currentPoint = currentPoint + (currentDir * nextDistance);
        }
        if (!didScatter) {
            currentPoint = entry;
            currentDir = -normalize(currentPoint);
        }
        const uint sampleCount = 8;
        vec3 tracedLight = vec3(0.0f);
        for (uint i = 0; i < sampleCount; i++) tracedLight += TracePathTrain(currentPoint, currentDir);
        tracedLight /= float(sampleCount);
        outPos = currentPoint; outDir = currentDir; outTarget = tracedLight;
    }
    // train.comp:1228-1259 for global invocation (x, y) [A18: no y flip, uv at the corner].
    void TrainInvocation(uint x, uint y, vec3& outPos, vec3& outDir, vec3& outTarget)
    {
        const vec2 fragUV = vec2(float(x) * S.WIDTH_FACTOR, float(y) * S.HEIGHT_FACTOR);
        const vec4 screenCoord = vec4((fragUV.x * 2.0f) - 1.0f, (fragUV.y * 2.0f) - 1.0f, 0.0f, 1.0f);
        const vec4 worldPos = S.camMat.invProjView * screenCoord;
        const vec3 pixelWorldPos = worldPos.xyz() / worldPos.w;
        preRand = S.volumeData.random.x * fragUV.x;
        prePreRand = S.volumeData.random.y * fragUV.y;
        const vec3 ro = S.camera.pos;
        vec3 rd = normalize(pixelWorldPos - ro);
        vec3 ee[2];
        find_entry_exit(ro, rd, ee);
        if (sky_sdf(ee[0]) > MAX_RAY_DISTANCE) rd = -normalize(ro);
        TracePathAndRecord(ro, rd, outPos, outDir, outTarget);
    }
};

// ---- optimiser steps ----
inline bool IsNanOrInf(float x) { return isnan(x) || isinf(x) || abs(x) > 1000.0f; }   // nrc-step:200-203, mrhe-step:27-30

// nrc-step.comp:205-392 over the packed arena.  Weights get the bad(w)->0 scrub, biases do
// not [A14]; "Adam" here is first-moment momentum only [A16].
inline void NrcStep(Scene& S)
{
    const float border = 1000.0f;
    const float beta1 = S.nrcConfig.beta1;
    float adamC1 = 1.0f, adamC2 = 1.0f;   // bias corrections 1/(1 - beta^t)
    if (S.adam) {
        if (S.secondMoment.size() != kParamCount) S.secondMoment.assign(kParamCount, 0.0f);
        S.nrcSteps++;
        adamC1 = 1.0f / (1.0f - pow(beta1, float(S.nrcSteps)));
        adamC2 = 1.0f / (1.0f - pow(S.adamBeta2, float(S.nrcSteps)));
    }
    for (size_t i = 0; i < kParamCount; i++) {
        const bool isWeight = i < kBiasOffset[0];
        float weight = S.params[i];
        float deltaWeight = S.deltaParams[i];
        if (IsNanOrInf(deltaWeight)) deltaWeight = border * sign(deltaWeight);   // [A13]
        deltaWeight += -weight * S.nrcConfig.weightDecay;
        float d = ((1.0f - beta1) * deltaWeight) + (beta1 * S.momentum[i]);
        S.momentum[i] = d;
        if (S.adam) {   // deltas are NEGATIVE gradients: w += lr * mhat / (sqrt(vhat) + eps)
            const float v = (S.adamBeta2 * S.secondMoment[i]) + ((1.0f - S.adamBeta2) * (deltaWeight * deltaWeight));
            S.secondMoment[i] = v;
            d = (d * adamC1) / (sqrt(v * adamC2) + S.adamEps);
        }
        S.params[i] = S.params[i] + d * S.nrcConfig.learningRate;
        if (isWeight && IsNanOrInf(S.params[i])) S.params[i] = 0.0f;
        S.deltaParams[i] = 0.0f;
    }
}

// mrhe-step.comp:32-74
inline void MrheStep(Scene& S)
{
    const float border = 100.0f;
    const size_t maxIndex = (size_t)S.mrhe.levelCount * S.mrhe.featureCount * S.mrhe.hashTableSize;
    float adamC1 = 1.0f, adamC2 = 1.0f;
    if (S.adam) {
        if (S.mrFirstMoment.size() != maxIndex) { S.mrFirstMoment.assign(maxIndex, 0.0f); S.mrSecondMoment.assign(maxIndex, 0.0f); }
        S.mrheSteps++;
        adamC1 = 1.0f / (1.0f - pow(S.mrheBeta1, float(S.mrheSteps)));
        adamC2 = 1.0f / (1.0f - pow(S.adamBeta2, float(S.mrheSteps)));
    }
    for (size_t index = 0; index < maxIndex; index++) {
        float deltaWeight = S.mrDeltaHashTable[index];
        const float weight = S.mrHashTable[index];
        if (IsNanOrInf(deltaWeight)) deltaWeight = border * sign(deltaWeight);
        deltaWeight += -weight * S.mrhe.weightDecay;
        if (S.adam) {
            const float m = (S.mrheBeta1 * S.mrFirstMoment[index]) + ((1.0f - S.mrheBeta1) * deltaWeight);
            const float v = (S.adamBeta2 * S.mrSecondMoment[index]) + ((1.0f - S.adamBeta2) * (deltaWeight * deltaWeight));
            S.mrFirstMoment[index] = m;
            S.mrSecondMoment[index] = v;
            deltaWeight = (m * adamC1) / (sqrt(v * adamC2) + S.adamEps);
        }
        S.mrHashTable[index] = S.mrHashTable[index] + deltaWeight * S.mrhe.learningRate;
        if (IsNanOrInf(S.mrHashTable[index])) S.mrHashTable[index] = 0.0f;
        S.mrDeltaHashTable[index] = 0.0f;
    }
}

}  // namespace oracle_impl
}  // namespace glsl
namespace oracle = glsl::oracle_impl;
