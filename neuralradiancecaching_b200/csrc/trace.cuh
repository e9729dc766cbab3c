// trace.cuh -- data layout of the wavefront volume tracer (shared by trace.cu and renderer.cu).
//
// The reference traces one path per shader invocation, start to finish, lighting included
// (nrc-forward.frag:757-834, nrc-train.comp:1122-1226).  Its RNG is a serial chain, but nothing the chain
// or the walk's control flow depends on comes from the expensive part -- the 96..224 density taps per lit
// vertex only produce radiance VALUES.  So the frame is cut into three kinds of kernels:
//
//   walk     one thread per pixel / per training sample: the serial part only (RNG, direction sampling,
//            box SDF marching, one density tap per step).  Emits a VERTEX record for every lit vertex.
//   light    one thread per vertex, every vertex the same amount of work: directional light, point light,
//            (render only) 8 environment samples, and the segment transmittance -- all the taps.
//   combine  one thread per pixel / sample: folds its vertices' results in path order
//            (L += T * density * light; T *= t_r), emits the NRC query or the training record.
//
// Vertices live in a pool filled through a warp-aggregated atomic counter and are chained per path.
#pragma once
#include "nrc_internal.cuh"

struct VertexPool {
    float4* a;        // pos.xyz, density
    float4* b;        // dir.xyz, next vertex (int bits; -1 = end of path)
    float4* c;        // lastPoint.xyz, unused
    float4* rnd;      // render only: [cap][4] = the 16 uniform draws of the 8 environment samples
    float4* out;      // light.rgb, t_r          (written by the light kernel)
    uint32_t* count;  // vertices allocated so far (device)
    uint32_t* overflow;
    uint32_t cap;
};

// Per render path (pixel * spp + s)
struct RenderPaths {
    int* head;            // first vertex or -1
    float4* qposPhase;    // terminal NRC query: pos.xyz, dirPhase   (valid iff qdir.w != 0)
    float4* qdirValid;    // dir.xyz, valid flag
    float4* result;       // L.rgb (non-NN part), T          (combine)
    float* qweight;       // T * dirPhase                     (combine)
    int* qslot;           // index into the compacted query list or -1
    // two-launch render walk (trace.cu): paths still alive after the head launch, compacted; 6 float4 planes of spillStride records
    float4* spill;
    uint32_t* spillCount;
    uint32_t spillStride;
};

struct QueryList {
    float* pos;        // [cap][3]
    float* dir;        // [cap][3]
    float* rgb;        // [cap][3]   (inference output)
    uint32_t* count;
    uint32_t cap;
};

struct TrainPaths {
    int* head;            // [nSamples][8]
    float* pos;           // [nSamples][3]  training record position
    float* dir;           // [nSamples][3]
    float* target;        // [nSamples][3]
};

// counters[.][4] = randDraws, litVertices, walkSteps, queryEmitted
// rows y0 + k*ystride (k < rows) of the frame; train invocations i0 + k*istride (k < n)
nrc_status launch_render_walk(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                              VertexPool pool, RenderPaths paths, float4* background, uint32_t* counters);
nrc_status launch_train_walk(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t i0, uint32_t n, uint32_t istride, VertexPool pool,
                             TrainPaths paths, uint32_t* counters);
nrc_status launch_light(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, VertexPool pool, bool render, unsigned long long* tapCounter, int kid);
nrc_status launch_render_combine(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                                 VertexPool pool, RenderPaths paths, QueryList q);
nrc_status launch_train_combine(nrc_ctx* ctx, cudaStream_t st, uint32_t nLocal, VertexPool pool, TrainPaths paths,
                                const unsigned long long* taps, unsigned long long* totals);
nrc_status launch_composite(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                            RenderPaths paths, QueryList q, const float4* background, float4* frame, uint32_t* frame8, float* queryProbe,
                            const uint32_t* scalars, const unsigned long long* taps, unsigned long long* totals);
nrc_status launch_volume_sample(nrc_volume* vol, const float* pos, size_t n, float* out);
nrc_status launch_build_bricks(nrc_volume* vol, const float* densityDev);
