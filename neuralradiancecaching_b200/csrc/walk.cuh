// walk.cuh -- the serial random walks of the tracer as host/device templates.
//
// The kernels in trace.cu instantiate these with an emitter that appends vertices to the device pool;
// tests/host_emul instantiates the very same code on the CPU build machine (no GPU there) with a recording
// emitter and compares every draw count, vertex and query with the oracle.  It is a debugging harness for the
// device code, not a product path: nothing in the library calls the host instantiation.
//
// Reference: nrc-forward.vert:28-40, nrc-forward.frag:757-870 (render); nrc-train.comp:1122-1259 (train).
#pragma once
#include "strict.cuh"

struct WalkParams {
    float invProjView[16];
    float camPos[3];
    float randomX, randomY;     // volumeData.random.xy
    float g;
    unsigned useNN, showNonNN;
    unsigned width, height;     // frame
    unsigned trainW;
    float widthFactor, heightFactor;
};

struct WalkCounters { unsigned lit, steps, sdf, emitted; };

struct WalkQuery { f3 pos, dir, lastDir; bool valid; };

// One walk step shared by all walks: exit along dir, next point (nrc-forward.frag:813-818).
NRC_HD f3 walk_advance(const Box& box, f3 cur, f3 dir, Rng& rng, unsigned& sdf)
{
    f3 en, ex;
    find_entry_exit(box, cur, dir, en, ex, sdf);
    float maxDistance = NRC_MUL(s_length(s_sub(ex, cur)), 0.1f);
    float nextDistance = NRC_MUL(rng_next(rng), maxDistance);
    return s_madd(cur, dir, nextDistance);
}

// Camera ray + RNG seed of render pixel (px, py): fragUV at the pixel centre, y flipped at the vertices (A18).
NRC_HD void render_ray(const WalkParams& P, unsigned px, unsigned py, f3& ro, f3& rd, Rng& rng)
{
    float u = NRC_DIV(NRC_ADD((float)px, 0.5f), (float)P.width);
    float v = NRC_DIV(NRC_ADD((float)py, 0.5f), (float)P.height);
    float sx = NRC_SUB(NRC_MUL(u, 2.0f), 1.0f);
    float sy = NRC_MUL(NRC_SUB(NRC_MUL(v, 2.0f), 1.0f), -1.0f);
    ro = mk3(P.camPos[0], P.camPos[1], P.camPos[2]);
    rd = s_normalize(s_sub(unproject(P.invProjView, sx, sy), ro));
    rng.pre = NRC_MUL(P.randomX, u);
    rng.prePre = NRC_MUL(P.randomY, v);
    rng.draws = 0;
}

// TracePath(ro, rd, useNN) of the render shader without lighting and Forward(): walks from `entry`, calls
// em.vertex(pos, density, dir, last, rands16) for every lit vertex (false = pool full, stop) and fills `q` when the
// path terminates into the cache.
//
// The loop is resumable: render_walk_steps runs steps [i0, i1) on a WalkState and says whether the path is still alive, so that
// the render kernel can stop after a few steps, compact the surviving paths and finish them in a second launch with full warps
// (trace.cu).  render_walk_path is the whole walk in one call (host emulation, spp > 1).
struct WalkState { f3 cur, last, dir, lastDir; float termProb; };

NRC_HD WalkState walk_state_begin(f3 entry, f3 rd)
{
    WalkState s;
    s.cur = entry; s.last = entry; s.dir = rd; s.lastDir = mk3(0.0f, 0.0f, 0.0f); s.termProb = 1.0f;
    return s;
}

#if defined(__CUDACC__)
#pragma nv_exec_check_disable
#endif
template <class Emitter>
NRC_HD bool render_walk_steps(const WalkParams& P, const Box& box, const Grid& grid, WalkState& s, int i0, int i1, Rng& rng, Emitter& em,
                              WalkCounters& wc, WalkQuery& q)
{
    const bool useNN = (P.useNN == 1u);
    for (int i = i0; i < i1; i++) {
        wc.steps++;
        float density = grid_density(grid, s.cur);
        if (density > 0.0f) {
            if (useNN) {
                if (rng_next(rng) > s.termProb) {
                    if (P.showNonNN == 0u) {
                        q.pos = s.cur; q.dir = s.dir; q.lastDir = s.lastDir; q.valid = true;   // A19: lastDir = 0 before the first scatter
                        wc.emitted++;
                    }
                    return false;
                }
                s.termProb = NRC_MUL(s.termProb, 0.5f);
            }
            float r[16];
            for (int k = 0; k < 16; k++) r[k] = rng_next(rng);   // the 8 NewRayDir calls of SampleHdrEnvMap (fwd.frag:717-725)
            if (!em.vertex(s.cur, density, s.dir, s.last, r)) return false;
            wc.lit++;
            s.last = s.cur;
            s.lastDir = s.dir;
            s.dir = new_ray_dir(s.dir, P.g, rng);
        }
        s.cur = walk_advance(box, s.cur, s.dir, rng, wc.sdf);
    }
    return true;
}

#define NRC_RENDER_WALK_STEPS 32

#if defined(__CUDACC__)
#pragma nv_exec_check_disable
#endif
template <class Emitter>
NRC_HD void render_walk_path(const WalkParams& P, const Box& box, const Grid& grid, f3 entry, f3 rd, Rng& rng, Emitter& em,
                             WalkCounters& wc, WalkQuery& q)
{
    WalkState s = walk_state_begin(entry, rd);
    q.valid = false;
    render_walk_steps(P, box, grid, s, 0, NRC_RENDER_WALK_STEPS, rng, em, wc, q);
}

// main() of nrc-train.comp (:1228-1259) + the vertex-picking walk of TracePathAndLearn (:1173-1215): returns the
// training record's position and direction.
NRC_HD void train_pick_vertex(const WalkParams& P, const Box& box, const Grid& grid, unsigned inv, Rng& rng, WalkCounters& wc,
                              f3& outPos, f3& outDir)
{
    unsigned x = inv % P.trainW, y = inv / P.trainW;
    float u = NRC_MUL((float)x, P.widthFactor);     // uv at the corner, no y flip (A18)
    float v = NRC_MUL((float)y, P.heightFactor);
    float sx = NRC_SUB(NRC_MUL(u, 2.0f), 1.0f);
    float sy = NRC_SUB(NRC_MUL(v, 2.0f), 1.0f);
    f3 ro = mk3(P.camPos[0], P.camPos[1], P.camPos[2]);
    f3 rd = s_normalize(s_sub(unproject(P.invProjView, sx, sy), ro));
    rng.pre = NRC_MUL(P.randomX, u);
    rng.prePre = NRC_MUL(P.randomY, v);
    rng.draws = 0;
    f3 entry, exitp;
    find_entry_exit(box, ro, rd, entry, exitp, wc.sdf);
    if (sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE) {
        rd = s_neg(s_normalize(ro));
        find_entry_exit(box, ro, rd, entry, exitp, wc.sdf);   // TracePathAndLearn recomputes it with the new direction
    }
    f3 cur = entry, dir = rd;
    float termProb = 1.0f;
    bool didScatter = false;
    for (int i = 0; i < 64; i++) {
        if (rng_next(rng) > termProb) break;
        termProb = NRC_MUL(termProb, 0.5f);
        wc.steps++;
        if (grid_density(grid, cur) > 0.0f) {
            didScatter = true;
            dir = new_ray_dir(dir, P.g, rng);
        }
        cur = walk_advance(box, cur, dir, rng, wc.sdf);
    }
    if (!didScatter) {
        cur = entry;
        dir = s_neg(s_normalize(cur));
    }
    outPos = cur;
    outDir = dir;
}

#if defined(__CUDACC__)
#pragma nv_exec_check_disable
#endif
// TracePath(rayOrigin, rayDir) of the train shader (:1122-1171) without lighting.
// Written for latency, not instruction count (the train walk is a serial chain on a handful of warps): the scatter
// direction is computed from the next two draws BEFORE the density tap has returned, so the tap's memory latency, the
// RNG chain and the rotation arithmetic overlap; which of the two continuations is taken is a select.  Same draws,
// same operations, same results as the straightforward order.
template <class Emitter>
NRC_HD void train_walk_path(const WalkParams& P, const Box& box, const Grid& grid, f3 startP, f3 startD, Rng& rng, Emitter& em,
                            WalkCounters& wc)
{
    f3 en, ex;
    find_entry_exit(box, startP, startD, en, ex, wc.sdf);
    f3 c = en, last = en, d = startD;
    for (int i = 0; i < 64; i++) {
        wc.steps++;
        float density = grid_density(grid, c);
        Rng spec = rng;                                  // speculative: the two draws of NewRayDir
        f3 nd = new_ray_dir(d, P.g, spec);
        if (density > 0.0f) {
            if (!em.vertex(c, density, d, last, nullptr)) return;
            wc.lit++;
            last = c;
            d = nd;
            rng = spec;
        }
        c = walk_advance(box, c, d, rng, wc.sdf);
    }
}

// The same walk with the env-map term of TraceScene switched on (opt-in; train.comp:1118 has it commented out): a lit vertex
// first draws the 16 numbers of SampleHdrEnvMap's 8 NewRayDir calls, then the two of its own NewRayDir -- the render walk's order.
template <class Emitter>
NRC_HD void train_walk_path_env(const WalkParams& P, const Box& box, const Grid& grid, f3 startP, f3 startD, Rng& rng, Emitter& em,
                                WalkCounters& wc)
{
    f3 en, ex;
    find_entry_exit(box, startP, startD, en, ex, wc.sdf);
    f3 c = en, last = en, d = startD;
    for (int i = 0; i < 64; i++) {
        wc.steps++;
        float density = grid_density(grid, c);
        if (density > 0.0f) {
            float r[16];
            for (int k = 0; k < 16; k++) r[k] = rng_next(rng);
            if (!em.vertex(c, density, d, last, r)) return;
            wc.lit++;
            last = c;
            d = new_ray_dir(d, P.g, rng);
        }
        c = walk_advance(box, c, d, rng, wc.sdf);
    }
}
