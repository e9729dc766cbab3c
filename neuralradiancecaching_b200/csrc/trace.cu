// trace.cu -- wavefront volume tracer kernels; layout and rationale in trace.cuh.
//
// walk kernels use strict.cuh (bit-exact control flow against the oracle); light/combine kernels use fast
// math: their results are radiance values only.
#include "trace.cuh"
#include "strict.cuh"

// ---- small device helpers ---------------------------------------------------------------------------------------

__device__ __forceinline__ Box make_box(const TraceParams& P)
{
    Box b;
    b.half = mk3(P.skySize[0] * 0.5f, P.skySize[1] * 0.5f, P.skySize[2] * 0.5f);   // exact halvings
    b.exitJump = P.exitJump;
    return b;
}

__device__ __forceinline__ Grid make_grid(const TraceParams& P)
{
    Grid g;
    g.nx = P.nx; g.ny = P.ny; g.nz = P.nz;
    g.bricks = P.bricks;
    g.skySize = mk3(P.skySize[0], P.skySize[1], P.skySize[2]);
    g.densityFactor = P.densityFactor;
    return g;
}

// Warp-aggregated slot allocation: one atomic per converged group of lanes.
__device__ __forceinline__ uint32_t alloc_slot(uint32_t* counter)
{
    unsigned active = __activemask();
    int leader = __ffs(active) - 1;
    unsigned lane = threadIdx.x & 31;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(active));
    base = __shfl_sync(active, base, leader);
    return base + __popc(active & ((1u << lane) - 1u));
}

// hg_phase_func (nrc-forward.frag:584-590), value only
__device__ __forceinline__ float hg_phase(float g, float cosTheta)
{
    float g2 = g * g;
    float x = 1.0f + g2 - (2.0f * g * cosTheta);
    return 0.5f * (1.0f - g2) / (x * sqrtf(x));
}

__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// SampleHdrEnvMap(dir, hpm) (nrc-forward.frag:690-708): bilinear, clamp-to-edge, RGBA32F
__device__ __forceinline__ f3 env_lookup(const TraceParams& P, f3 d, float strength)
{
    float u = atan2f(d.z, d.x) * 0.1591f + 0.5f;
    float v = asinf(fminf(fmaxf(d.y, -1.0f), 1.0f)) * 0.3183f + 0.5f;
    float x = u * (float)P.envW - 0.5f, y = v * (float)P.envH - 0.5f;
    if (!(x == x) || !(y == y)) return mk3(0.0f, 0.0f, 0.0f);
    x = fminf(fmaxf(x, -1.0f), (float)P.envW);
    y = fminf(fmaxf(y, -1.0f), (float)P.envH);
    float fx = floorf(x), fy = floorf(y);
    float a = x - fx, b = y - fy;
    int i0 = (int)fx, j0 = (int)fy;
    int i1 = min(max(i0 + 1, 0), P.envW - 1), j1 = min(max(j0 + 1, 0), P.envH - 1);
    i0 = min(max(i0, 0), P.envW - 1); j0 = min(max(j0, 0), P.envH - 1);
    float4 t00 = __ldg(P.env + (size_t)j0 * P.envW + i0), t10 = __ldg(P.env + (size_t)j0 * P.envW + i1);
    float4 t01 = __ldg(P.env + (size_t)j1 * P.envW + i0), t11 = __ldg(P.env + (size_t)j1 * P.envW + i1);
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return mk3((w00 * t00.x + w10 * t10.x + w01 * t01.x + w11 * t11.x) * strength,
               (w00 * t00.y + w10 * t10.y + w01 * t01.y + w11 * t11.y) * strength,
               (w00 * t00.z + w10 * t10.z + w01 * t01.z + w11 * t11.z) * strength);
}

// One walk step shared by all walks: exit along dir, next point (nrc-forward.frag:813-818).
__device__ __forceinline__ f3 walk_advance(const Box& box, f3 cur, f3 dir, Rng& rng, unsigned& sdf)
{
    f3 en, ex;
    find_entry_exit(box, cur, dir, en, ex, sdf);
    float maxDistance = NRC_MUL(s_length(s_sub(ex, cur)), 0.1f);
    float nextDistance = NRC_MUL(rng_next(rng), maxDistance);
    return s_madd(cur, dir, nextDistance);
}

__device__ __forceinline__ void store_vertex(const VertexPool& pool, uint32_t slot, f3 pos, float density, f3 dir, f3 last)
{
    pool.a[slot] = make_float4(pos.x, pos.y, pos.z, density);
    pool.b[slot] = make_float4(dir.x, dir.y, dir.z, __int_as_float(-1));
    pool.c[slot] = make_float4(last.x, last.y, last.z, 0.0f);
}

// ---- render walk (nrc-forward.vert:28-40, nrc-forward.frag:757-870 minus lighting and Forward()) ------------------
__global__ void __launch_bounds__(128) k_render_walk(const __grid_constant__ TraceParams P, uint32_t y0, uint32_t y1, uint32_t spp,
                                                     VertexPool pool, RenderPaths paths, float4* __restrict__ background,
                                                     uint32_t* __restrict__ counters)
{
    uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rows = y1 - y0;
    if (local >= rows * P.width) return;
    uint32_t px = local % P.width, py = y0 + local / P.width;
    size_t pix = (size_t)py * P.width + px;
    const Box box = make_box(P);
    const Grid grid = make_grid(P);

    float u = NRC_DIV(NRC_ADD((float)px, 0.5f), (float)P.width);
    float v = NRC_DIV(NRC_ADD((float)py, 0.5f), (float)P.height);
    float sx = NRC_SUB(NRC_MUL(u, 2.0f), 1.0f);
    float sy = NRC_MUL(NRC_SUB(NRC_MUL(v, 2.0f), 1.0f), -1.0f);   // y flipped at the vertices (A18)
    f3 ro = mk3(P.camPos[0], P.camPos[1], P.camPos[2]);
    f3 rd = s_normalize(s_sub(unproject(P.invProjView, sx, sy), ro));
    Rng rng;
    rng.pre = NRC_MUL(P.random[0], u);
    rng.prePre = NRC_MUL(P.random[1], v);
    rng.draws = 0;
    unsigned sdf = 0, lit = 0, steps = 0, emitted = 0;

    f3 entry, exitp;
    find_entry_exit(box, ro, rd, entry, exitp, sdf);
    f3 envc = env_lookup(P, rd, P.envDirect);
    bool miss = sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE;
    background[pix] = make_float4(envc.x, envc.y, envc.z, miss ? 1.0f : 0.0f);
    const bool useNN = (P.useNN == 1u);

    for (uint32_t s = 0; s < spp; s++) {
        size_t path = pix * spp + s;
        int head = -1;
        int prevSlot = -1;
        float4 qp = make_float4(0.0f, 0.0f, 0.0f, 0.0f), qd = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (!miss) {
            f3 cur = entry, last = entry, dir = rd, lastDir = mk3(0.0f, 0.0f, 0.0f);
            float termProb = 1.0f;
            for (int i = 0; i < 32; i++) {
                steps++;
                float density = grid_density(grid, cur);
                if (density > 0.0f) {
                    if (useNN) {
                        if (rng_next(rng) > termProb) {
                            if (P.showNonNN == 0u) {
                                float ph = hg_phase(P.g, -dot3(dir, lastDir));   // A19: lastDir = 0 before the first scatter
                                qp = make_float4(cur.x, cur.y, cur.z, ph);
                                qd = make_float4(dir.x, dir.y, dir.z, 1.0f);
                                emitted++;
                            }
                            break;
                        }
                        termProb = NRC_MUL(termProb, 0.5f);
                    }
                    uint32_t slot = alloc_slot(pool.count);
                    if (slot >= pool.cap) { *pool.overflow = 1u; break; }
                    store_vertex(pool, slot, cur, density, dir, last);
                    float r[16];
#pragma unroll
                    for (int k = 0; k < 16; k++) r[k] = rng_next(rng);   // the 8 NewRayDir calls of SampleHdrEnvMap (fwd.frag:717-725)
#pragma unroll
                    for (int k = 0; k < 4; k++) pool.rnd[(size_t)slot * 4 + k] = make_float4(r[4 * k], r[4 * k + 1], r[4 * k + 2], r[4 * k + 3]);
                    if (prevSlot >= 0) pool.b[prevSlot].w = __int_as_float((int)slot); else head = (int)slot;
                    prevSlot = (int)slot;
                    lit++;
                    last = cur;
                    lastDir = dir;
                    dir = new_ray_dir(dir, P.g, rng);
                }
                cur = walk_advance(box, cur, dir, rng, sdf);
            }
        }
        paths.head[path] = head;
        paths.qposPhase[path] = qp;
        paths.qdirValid[path] = qd;
    }
    if (counters) {
        uint4 c = make_uint4(rng.draws, lit, steps, emitted);
        reinterpret_cast<uint4*>(counters)[pix] = c;
    }
}

// ---- train walk (nrc-train.comp:1122-1259 minus lighting and Backprop) ------------------------------------------------
__global__ void __launch_bounds__(64) k_train_walk(const __grid_constant__ TraceParams P, uint32_t i0, uint32_t i1, VertexPool pool,
                                                   TrainPaths paths, uint32_t* __restrict__ counters)
{
    uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= i1 - i0) return;
    uint32_t inv = i0 + local;
    uint32_t x = inv % P.trainW, y = inv / P.trainW;
    const Box box = make_box(P);
    const Grid grid = make_grid(P);

    float u = NRC_MUL((float)x, P.widthFactor);     // uv at the corner, no y flip (A18)
    float v = NRC_MUL((float)y, P.heightFactor);
    float sx = NRC_SUB(NRC_MUL(u, 2.0f), 1.0f);
    float sy = NRC_SUB(NRC_MUL(v, 2.0f), 1.0f);
    f3 ro = mk3(P.camPos[0], P.camPos[1], P.camPos[2]);
    f3 rd = s_normalize(s_sub(unproject(P.invProjView, sx, sy), ro));
    Rng rng;
    rng.pre = NRC_MUL(P.random[0], u);
    rng.prePre = NRC_MUL(P.random[1], v);
    rng.draws = 0;
    unsigned sdf = 0, lit = 0, steps = 0;

    f3 entry, exitp;
    find_entry_exit(box, ro, rd, entry, exitp, sdf);
    if (sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE) {
        rd = s_neg(s_normalize(ro));
        find_entry_exit(box, ro, rd, entry, exitp, sdf);   // TracePathAndLearn recomputes it with the new direction
    }
    // TracePathAndLearn (:1173-1215): pick the training vertex
    f3 cur = entry, dir = rd;
    {
        float termProb = 1.0f;
        bool didScatter = false;
        for (int i = 0; i < 64; i++) {
            if (rng_next(rng) > termProb) break;
            termProb = NRC_MUL(termProb, 0.5f);
            steps++;
            if (grid_density(grid, cur) > 0.0f) {
                didScatter = true;
                dir = new_ray_dir(dir, P.g, rng);
            }
            cur = walk_advance(box, cur, dir, rng, sdf);
        }
        if (!didScatter) {
            cur = entry;
            dir = s_neg(s_normalize(cur));
        }
    }
    paths.pos[3 * (size_t)local] = cur.x; paths.pos[3 * (size_t)local + 1] = cur.y; paths.pos[3 * (size_t)local + 2] = cur.z;
    paths.dir[3 * (size_t)local] = dir.x; paths.dir[3 * (size_t)local + 1] = dir.y; paths.dir[3 * (size_t)local + 2] = dir.z;
    // 8 x TracePath (:1122-1171), RNG chained from one to the next
    const f3 startP = cur, startD = dir;
    for (int p = 0; p < 8; p++) {
        f3 en, ex;
        find_entry_exit(box, startP, startD, en, ex, sdf);
        f3 c = en, last = en, d = startD;
        int head = -1, prevSlot = -1;
        for (int i = 0; i < 64; i++) {
            steps++;
            float density = grid_density(grid, c);
            if (density > 0.0f) {
                uint32_t slot = alloc_slot(pool.count);
                if (slot >= pool.cap) { *pool.overflow = 1u; break; }
                store_vertex(pool, slot, c, density, d, last);
                if (prevSlot >= 0) pool.b[prevSlot].w = __int_as_float((int)slot); else head = (int)slot;
                prevSlot = (int)slot;
                lit++;
                last = c;
                d = new_ray_dir(d, P.g, rng);
            }
            c = walk_advance(box, c, d, rng, sdf);
        }
        paths.head[(size_t)local * 8 + p] = head;
    }
    if (counters) reinterpret_cast<uint4*>(counters)[local] = make_uint4(rng.draws, lit, steps, 0u);
}
