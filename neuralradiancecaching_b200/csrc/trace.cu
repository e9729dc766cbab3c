// trace.cu -- wavefront volume tracer kernels; layout and rationale in trace.cuh.
//
// walk kernels use strict.cuh (bit-exact control flow against the oracle); light/combine kernels use fast
// math: their results are radiance values only.
#include "trace.cuh"
#include "walk.cuh"
#include <stdlib.h>

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
    g.invSize = mk3(__frcp_rn(g.skySize.x), __frcp_rn(g.skySize.y), __frcp_rn(g.skySize.z));
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

// texture(sampler, u) of a single-channel float image: linear, clamp-to-edge (the two inverse-CDF images, HdrEnvMap.cpp:126-148)
__device__ __forceinline__ float tex1_clamp(const float* __restrict__ row, int w, float uc)
{
    float u = uc * (float)w - 0.5f;
    if (!(u == u)) return 0.0f;
    u = fminf(fmaxf(u, -1.0f), (float)w);
    float fu = floorf(u), a = u - fu;
    int i0 = (int)fu, i1 = i0 + 1;
    i0 = min(max(i0, 0), w - 1); i1 = min(max(i1, 0), w - 1);
    return (1.0f - a) * __ldg(row + i0) + a * __ldg(row + i1);
}

__device__ __forceinline__ float tex2_clamp(const float* __restrict__ img, int w, int h, float uc, float vc)
{
    float x = uc * (float)w - 0.5f, y = vc * (float)h - 0.5f;
    if (!(x == x) || !(y == y)) return 0.0f;
    x = fminf(fmaxf(x, -1.0f), (float)w);
    y = fminf(fmaxf(y, -1.0f), (float)h);
    float fx = floorf(x), fy = floorf(y), a = x - fx, b = y - fy;
    int i0 = (int)fx, j0 = (int)fy;
    int i1 = min(max(i0 + 1, 0), w - 1), j1 = min(max(j0 + 1, 0), h - 1);
    i0 = min(max(i0, 0), w - 1); j0 = min(max(j0, 0), h - 1);
    float t00 = __ldg(img + (size_t)j0 * w + i0), t10 = __ldg(img + (size_t)j0 * w + i1);
    float t01 = __ldg(img + (size_t)j1 * w + i0), t11 = __ldg(img + (size_t)j1 * w + i1);
    return ((1.0f - a) * (1.0f - b) * t00 + a * (1.0f - b) * t10) + ((1.0f - a) * b * t01 + a * b * t11);
}

// texture(hdrEnvMap, uv).xyz * strength
__device__ __forceinline__ f3 env_texel(const TraceParams& P, float u, float v, float strength)
{
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

__device__ __forceinline__ WalkParams make_walk(const TraceParams& P)
{
    WalkParams w;
#pragma unroll
    for (int i = 0; i < 16; i++) w.invProjView[i] = P.invProjView[i];
    w.camPos[0] = P.camPos[0]; w.camPos[1] = P.camPos[1]; w.camPos[2] = P.camPos[2];
    w.randomX = P.random[0]; w.randomY = P.random[1];
    w.g = P.g; w.useNN = P.useNN; w.showNonNN = P.showNonNN;
    w.width = P.width; w.height = P.height; w.trainW = P.trainW;
    w.widthFactor = P.widthFactor; w.heightFactor = P.heightFactor;
    return w;
}

// Appends lit vertices to the pool and chains them per path.
struct PoolEmitter {
    VertexPool pool;
    int head, prev;
    __device__ __forceinline__ void begin() { head = -1; prev = -1; }
    __device__ __forceinline__ bool vertex(f3 pos, float density, f3 dir, f3 last, const float* r)
    {
        uint32_t slot = alloc_slot(pool.count);
        if (slot >= pool.cap) { *pool.overflow = 1u; return false; }
        pool.a[slot] = make_float4(pos.x, pos.y, pos.z, density);
        pool.b[slot] = make_float4(dir.x, dir.y, dir.z, __int_as_float(-1));
        pool.c[slot] = make_float4(last.x, last.y, last.z, 0.0f);
        if (r) {
#pragma unroll
            for (int k = 0; k < 4; k++) pool.rnd[(size_t)slot * 4 + k] = make_float4(r[4 * k], r[4 * k + 1], r[4 * k + 2], r[4 * k + 3]);
        }
        if (prev >= 0) pool.b[prev].w = __int_as_float((int)slot); else head = (int)slot;
        prev = (int)slot;
        return true;
    }
};

// Emitter of the train walk.  There the walk is a latency-bound serial chain, and a slot allocation per vertex (an L2
// atomic whose result the next instructions need) would sit on it.  Each thread instead owns a chunk of TRAIN_CHUNK
// slots and fetches the next chunk one chunk ahead, so no atomic result is ever waited for.  Slots left unused at the
// end are marked dead (density < 0): the light kernel skips them and no path links to them.
#define TRAIN_CHUNK 16u
struct ChunkEmitter {
    VertexPool pool;
    int head, prev;
    uint32_t cur, end, next;
    __device__ __forceinline__ void init() { cur = 0u; end = 0u; next = atomicAdd(pool.count, TRAIN_CHUNK); }
    __device__ __forceinline__ void begin() { head = -1; prev = -1; }
    __device__ __forceinline__ bool vertex(f3 pos, float density, f3 dir, f3 last, const float* r)
    {
        if (cur == end) {
            cur = next;
            end = cur + TRAIN_CHUNK;
            next = atomicAdd(pool.count, TRAIN_CHUNK);
        }
        uint32_t slot = cur++;
        if (slot >= pool.cap) { *pool.overflow = 1u; return false; }
        pool.a[slot] = make_float4(pos.x, pos.y, pos.z, density);
        pool.b[slot] = make_float4(dir.x, dir.y, dir.z, __int_as_float(-1));
        pool.c[slot] = make_float4(last.x, last.y, last.z, 0.0f);
        if (r != nullptr)   // env-map term in the training targets (opt-in): the 16 draws of SampleHdrEnvMap, as in the render pool
            for (int k = 0; k < 4; k++) pool.rnd[(size_t)slot * 4 + k] = make_float4(r[4 * k], r[4 * k + 1], r[4 * k + 2], r[4 * k + 3]);
        if (prev >= 0) pool.b[prev].w = __int_as_float((int)slot); else head = (int)slot;
        prev = (int)slot;
        return true;
    }
    __device__ __forceinline__ void finish()
    {
        for (uint32_t s = cur; s < end; s++) if (s < pool.cap) pool.a[s].w = -1.0f;
        for (uint32_t s = next; s < next + TRAIN_CHUNK; s++) if (s < pool.cap) pool.a[s].w = -1.0f;
    }
};

// ---- render walk (nrc-forward.vert:28-40, nrc-forward.frag:757-870 minus lighting and Forward()) ------------------
// Two launches when spp == 1 and splitAt > 0.  On the bench frame 24 % of the paths end within 12 steps (they terminate into the cache)
// and 76 % run all 32 (empty space inside the box: a step never covers more than a tenth of the way out), so with one pixel per lane
// for the whole walk a warp carries 15 % dead lanes through the sphere traces (scripts/walk_divergence.py; ncu: 27 of 32 lanes on
// the march lines).  The head launch walks `splitAt` steps and appends the state of every path that is still alive to a compacted list
// (80 bytes per path, one warp-aggregated atomic); the tail launch finishes those with full warps.  The walk itself is the same code
// on the same state (walk.cuh: render_walk_steps), so every draw and every vertex is what the single launch produces; only the
// order of the vertices in the pool differs, and nothing reads that order.
__device__ __forceinline__ void render_path_store(const TraceParams& P, RenderPaths paths, size_t path, int head, const WalkQuery& q)
{
    float4 qp = make_float4(0.0f, 0.0f, 0.0f, 0.0f), qd = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (q.valid) {
        float ph = hg_phase(P.g, -dot3(q.dir, q.lastDir));   // dirPhase (fwd.frag:785)
        qp = make_float4(q.pos.x, q.pos.y, q.pos.z, ph);
        qd = make_float4(q.dir.x, q.dir.y, q.dir.z, 1.0f);
    }
    paths.head[path] = head;
    paths.qposPhase[path] = qp;
    paths.qdirValid[path] = qd;
}

template <int MB>
__global__ void __launch_bounds__(128, MB) k_render_walk(const __grid_constant__ TraceParams P, uint32_t y0, uint32_t rows, uint32_t ystride,
                                                     uint32_t spp, VertexPool pool, RenderPaths paths, float4* __restrict__ background,
                                                     uint32_t* __restrict__ counters, int splitAt)
{
    // A warp = an 8 x 4 tile of this rank's pixels (rows of the partition), not 32 pixels of one row: neighbouring paths in BOTH directions
    // take similar walks, so fewer lanes idle through the marches (-0.5 % on the frame against the row-major mapping).  Which lane walks
    // which pixel changes nothing a pixel computes.
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t tilesX = (P.width + 7u) >> 3, tile = local >> 5, lane = local & 31u;
    const uint32_t px = (tile % tilesX) * 8u + (lane & 7u);
    const uint32_t prow = (tile / tilesX) * 4u + (lane >> 3);
    if (px >= P.width || prow >= rows) return;
    uint32_t py = y0 + prow * ystride;
    size_t pix = (size_t)py * P.width + px;
    const Box box = make_box(P);
    const Grid grid = make_grid(P);
    const WalkParams W = make_walk(P);
    f3 ro, rd;
    Rng rng;
    render_ray(W, px, py, ro, rd, rng);
    WalkCounters wc = {0u, 0u, 0u, 0u};
    f3 entry, exitp;
    find_entry_exit(box, ro, rd, entry, exitp, wc.sdf);
    f3 envc = env_lookup(P, rd, P.envDirect);
    bool miss = sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE;
    background[pix] = make_float4(envc.x, envc.y, envc.z, miss ? 1.0f : 0.0f);
    PoolEmitter em;
    em.pool = pool;
    const int stepsHere = splitAt > 0 ? splitAt : NRC_RENDER_WALK_STEPS;   // splitAt > 0 only with spp == 1 (launch_render_walk)
    for (uint32_t s = 0; s < spp; s++) {
        size_t path = pix * spp + s;
        em.begin();
        WalkQuery q;
        q.valid = false;
        if (!miss) {
            WalkState ws = walk_state_begin(entry, rd);
            bool alive = render_walk_steps(W, box, grid, ws, 0, stepsHere, rng, em, wc, q);
            if (alive && splitAt > 0) {
                const unsigned m = __activemask(), lane = threadIdx.x & 31u;
                const int leader = __ffs(m) - 1;
                unsigned base = 0u;
                if ((int)lane == leader) base = atomicAdd(paths.spillCount, (unsigned)__popc(m));
                const size_t slot = __shfl_sync(m, base, leader) + __popc(m & ((1u << lane) - 1u));
                const size_t n = paths.spillStride;
                paths.spill[slot] = make_float4(ws.cur.x, ws.cur.y, ws.cur.z, ws.termProb);
                paths.spill[n + slot] = make_float4(ws.last.x, ws.last.y, ws.last.z, rng.pre);
                paths.spill[2 * n + slot] = make_float4(ws.dir.x, ws.dir.y, ws.dir.z, rng.prePre);
                paths.spill[3 * n + slot] = make_float4(ws.lastDir.x, ws.lastDir.y, ws.lastDir.z, __uint_as_float((unsigned)pix));
                paths.spill[4 * n + slot] = make_float4(__int_as_float(em.head), __int_as_float(em.prev), __uint_as_float(rng.draws),
                                                        __uint_as_float(wc.lit));
                return;   // the tail launch writes this path's outputs
            }
        }
        render_path_store(P, paths, path, em.head, q);
    }
    if (counters) reinterpret_cast<uint4*>(counters)[pix] = make_uint4(rng.draws, wc.lit, wc.steps, wc.emitted);
}

template <int MB>
__global__ void __launch_bounds__(128, MB) k_render_walk_tail(const __grid_constant__ TraceParams P, VertexPool pool, RenderPaths paths,
                                                          uint32_t* __restrict__ counters, int splitAt)
{
    const size_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= *paths.spillCount) return;
    const size_t n = paths.spillStride;
    const float4 a = paths.spill[t], b = paths.spill[n + t], c = paths.spill[2 * n + t], d = paths.spill[3 * n + t], e = paths.spill[4 * n + t];
    const Box box = make_box(P);
    const Grid grid = make_grid(P);
    const WalkParams W = make_walk(P);
    WalkState ws;
    ws.cur = mk3(a.x, a.y, a.z); ws.termProb = a.w;
    ws.last = mk3(b.x, b.y, b.z);
    ws.dir = mk3(c.x, c.y, c.z);
    ws.lastDir = mk3(d.x, d.y, d.z);
    Rng rng;
    rng.pre = b.w; rng.prePre = c.w; rng.draws = __float_as_uint(e.z);
    const size_t pix = __float_as_uint(d.w);
    PoolEmitter em;
    em.pool = pool;
    em.head = __float_as_int(e.x); em.prev = __float_as_int(e.y);
    WalkCounters wc = {__float_as_uint(e.w), (unsigned)splitAt, 0u, 0u};
    WalkQuery q;
    q.valid = false;
    render_walk_steps(W, box, grid, ws, splitAt, NRC_RENDER_WALK_STEPS, rng, em, wc, q);
    render_path_store(P, paths, pix, em.head, q);
    if (counters) reinterpret_cast<uint4*>(counters)[pix] = make_uint4(rng.draws, wc.lit, wc.steps, wc.emitted);
}

// ---- train walk (nrc-train.comp:1122-1259 minus lighting and Backprop) ------------------------------------------------
// A training sample is ONE long serial chain (1 + 8 walks of 64 steps, the RNG chained through all of them) and a frame
// has only ~10^4 of them: ~300 warps, latency-bound, a few per cent of the machine's instructions -- but the longest chain of
// the frame.  It runs on a stream of its own, one frame ahead of the weights (renderer.cu), so only the next train walk waits
// for it, and consecutive frames' walks may be in flight together (renderer.cu).  Reserving SMs for it made the walk faster and the
// frame slower (code removed).  The launch bound keeps the kernel at 64 registers.
template <bool ENV>
__global__ void __launch_bounds__(1024, 1) k_train_walk(const __grid_constant__ TraceParams P, uint32_t i0, uint32_t n, uint32_t istride,
                                                        VertexPool pool, TrainPaths paths, uint32_t* __restrict__ counters)
{
    // 32 samples per warp: fewer (more, emptier warps) lose -- the walk's time is instruction latency per warp, not divergence, and the
    // extra warps compete with the render walk (profiles/r02_walks_in_flight.md)
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsPerBlock = blockDim.x >> 5;
    const uint32_t totalWarps = (n + 31u) / 32u;
    const Box box = make_box(P);
    const Grid grid = make_grid(P);
    const WalkParams W = make_walk(P);
    for (uint32_t w = blockIdx.x * warpsPerBlock + (threadIdx.x >> 5); w < totalWarps; w += gridDim.x * warpsPerBlock) {
    const uint32_t local = w * 32u + lane;
    if (local >= n) continue;
    ChunkEmitter em;
    em.pool = pool;
    em.init();                      // the first chunk arrives while the vertex-picking walk runs
    Rng rng;
    WalkCounters wc = {0u, 0u, 0u, 0u};
    f3 cur, dir;
    train_pick_vertex(W, box, grid, i0 + local * istride, rng, wc, cur, dir);
    paths.pos[3 * (size_t)local] = cur.x; paths.pos[3 * (size_t)local + 1] = cur.y; paths.pos[3 * (size_t)local + 2] = cur.z;
    paths.dir[3 * (size_t)local] = dir.x; paths.dir[3 * (size_t)local + 1] = dir.y; paths.dir[3 * (size_t)local + 2] = dir.z;
    for (int p = 0; p < 8; p++) {   // 8 x TracePath (:1217-1222), RNG chained from one to the next
        em.begin();
        if (ENV) train_walk_path_env(W, box, grid, cur, dir, rng, em, wc);
        else train_walk_path(W, box, grid, cur, dir, rng, em, wc);
        paths.head[(size_t)local * 8 + p] = em.head;
    }
    em.finish();
    if (counters) reinterpret_cast<uint4*>(counters)[local] = make_uint4(rng.draws, wc.lit, wc.steps, 0u);
    }
}

// ---- light kernel: every density tap of the frame happens here -----------------------------------------------------
// A tap is ~45 instructions: the segment is stepped in TEXEL space (u_i = u0 + i*du: one FFMA per axis), floor by
// FRND, the integer cell index read off the mantissa of (floor + 1.5*2^23) so that no F2I/I2F is needed, one 8-byte
// brick load, byte -> float by PRMT into the mantissa of 2^23 (the x-lerp differences are taken on the biased values,
// which is exact), 7 lerps.  The in-volume test is done once per SEGMENT: the region where the filter footprint
// exists (u in (-1, n) per axis) is convex, so a segment whose end points are inside never leaves it.  Values only
// (never steers control flow).
struct TapCtx {
    float sx, sy, sz, ox, oy, oz;   // texel coordinate = world * s + o  (= uvw * n - 0.5)
    float nx, ny, nz;
    int strideY, strideZ, bias;     // cell = iz*strideZ + iy*strideY + ix + bias on the magic-biased integers
    const uint2* bricks;
};

#define TAP_MAGIC 12582912.0f   // 1.5 * 2^23: float bits of (k + TAP_MAGIC) = 0x4B400000 + k for integer |k| < 2^22

__device__ __forceinline__ TapCtx make_tap(const TraceParams& P)
{
    TapCtx t;
    t.sx = (float)P.nx / P.skySize[0]; t.sy = (float)P.ny / P.skySize[1]; t.sz = (float)P.nz / P.skySize[2];
    t.ox = 0.5f * (float)P.nx - 0.5f; t.oy = 0.5f * (float)P.ny - 0.5f; t.oz = 0.5f * (float)P.nz - 0.5f;
    t.nx = (float)P.nx; t.ny = (float)P.ny; t.nz = (float)P.nz;
    t.strideY = P.nx + 1; t.strideZ = (P.nx + 1) * (P.ny + 1);
    // brick cell of texel-floor (i,j,k) is (i+1, j+1, k+1); the three indices arrive biased by 0x4B400000 each
    t.bias = (int)((1u + (unsigned)t.strideY + (unsigned)t.strideZ) * (1u - 0x4B400000u));
    t.bricks = P.bricks;
    return t;
}

// A tap in two halves so that a batch of taps can have all its brick loads in flight before the first blend:
// tap_cell  : texel-space position (u, v, s), -1 < u < nx etc. guaranteed by the caller -> brick index + the three fractions
// tap_blend : brick + fractions -> texel sum (0..255 scale)
__device__ __forceinline__ int tap_cell(const TapCtx& t, float u, float v, float s, float& a, float& b, float& c)
{
    float fu = floorf(u), fv = floorf(v), fs = floorf(s);
    a = u - fu; b = v - fv; c = s - fs;
    int ix = __float_as_int(fu + TAP_MAGIC), iy = __float_as_int(fv + TAP_MAGIC), iz = __float_as_int(fs + TAP_MAGIC);
    return iz * t.strideZ + iy * t.strideY + ix + t.bias;
}

__device__ __forceinline__ float tap_blend(uint2 q, float a, float b, float c)
{
    // byte k of the word -> float bits 0x4B0000kk = 2^23 + k
    float b000 = __uint_as_float(__byte_perm(q.x, 0x4B000000u, 0x7650)), b100 = __uint_as_float(__byte_perm(q.x, 0x4B000000u, 0x7651));
    float b010 = __uint_as_float(__byte_perm(q.x, 0x4B000000u, 0x7652)), b110 = __uint_as_float(__byte_perm(q.x, 0x4B000000u, 0x7653));
    float b001 = __uint_as_float(__byte_perm(q.y, 0x4B000000u, 0x7650)), b101 = __uint_as_float(__byte_perm(q.y, 0x4B000000u, 0x7651));
    float b011 = __uint_as_float(__byte_perm(q.y, 0x4B000000u, 0x7652)), b111 = __uint_as_float(__byte_perm(q.y, 0x4B000000u, 0x7653));
    float x00 = fmaf(a, b100 - b000, b000 - 8388608.0f), x10 = fmaf(a, b110 - b010, b010 - 8388608.0f);
    float x01 = fmaf(a, b101 - b001, b001 - 8388608.0f), x11 = fmaf(a, b111 - b011, b011 - 8388608.0f);
    float y0 = fmaf(b, x10 - x00, x00), y1 = fmaf(b, x11 - x01, x01);
    return fmaf(c, y1 - y0, y0);
}

__device__ __forceinline__ float tap_texel(const TapCtx& t, float u, float v, float s)
{
    float a, b, c;
    int cell = tap_cell(t, u, v, s, a, b, c);
    return tap_blend(__ldg(t.bricks + cell), a, b, c);
}

#define TAP_BATCH 8   // brick loads in flight per thread (ncu r01f: one load at a time left the light kernels 35-57 % long-scoreboard)

// ---- two taps per instruction: Blackwell's packed fp32 pair arithmetic (add/sub/fma .f32x2 -> FADD2 / FFMA2) -------------------
// The light kernels are bound by instruction issue (profiles/r01k: 65-74 % issue-active), and a tap is 46 instructions of which 31 are
// fp32 adds and FMAs.  Taps i and i+1 of a segment run the same arithmetic on different data, so they go through the pipe as the two
// halves of 64-bit register pairs: positions, fractions, the magic-biased cell coordinates and all 7 lerps are one packed instruction
// per PAIR of taps -- 31 instructions per tap instead of 46.  Every lane computes exactly what the scalar code computed (IEEE
// binary32, same operation order); only the order of the final sum changes (even and odd taps are summed separately).
typedef unsigned long long f2pair;
__device__ __forceinline__ f2pair pk2(float lo, float hi) { f2pair r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(f2pair v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2pair add2(f2pair a, f2pair b) { f2pair r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2pair sub2(f2pair a, f2pair b) { f2pair r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2pair fma2(f2pair a, f2pair b, f2pair c) { f2pair r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// texel-space positions of two taps -> their brick indices and the three fraction pairs
__device__ __forceinline__ void tap_cell2(const TapCtx& t, f2pair u, f2pair v, f2pair s, int& cell0, int& cell1, f2pair& a, f2pair& b, f2pair& c)
{
    float ul, uh, vl, vh, sl, sh;
    upk2(u, ul, uh); upk2(v, vl, vh); upk2(s, sl, sh);
    const f2pair fu = pk2(floorf(ul), floorf(uh)), fv = pk2(floorf(vl), floorf(vh)), fs = pk2(floorf(sl), floorf(sh));
    a = sub2(u, fu); b = sub2(v, fv); c = sub2(s, fs);
    const f2pair magic = pk2(TAP_MAGIC, TAP_MAGIC);
    float xl, xh, yl, yh, zl, zh;
    upk2(add2(fu, magic), xl, xh); upk2(add2(fv, magic), yl, yh); upk2(add2(fs, magic), zl, zh);
    cell0 = __float_as_int(zl) * t.strideZ + __float_as_int(yl) * t.strideY + __float_as_int(xl) + t.bias;
    cell1 = __float_as_int(zh) * t.strideZ + __float_as_int(yh) * t.strideY + __float_as_int(xh) + t.bias;
}

// two bricks + fraction pairs -> the two texel sums (0..255 scale), lane for lane the arithmetic of tap_blend
__device__ __forceinline__ f2pair tap_blend2(uint2 q0, uint2 q1, f2pair a, f2pair b, f2pair c)
{
    const unsigned K = 0x4B000000u;
#define NRC_B2(w, sel) pk2(__uint_as_float(__byte_perm(q0.w, K, sel)), __uint_as_float(__byte_perm(q1.w, K, sel)))
    const f2pair b000 = NRC_B2(x, 0x7650), b100 = NRC_B2(x, 0x7651), b010 = NRC_B2(x, 0x7652), b110 = NRC_B2(x, 0x7653);
    const f2pair b001 = NRC_B2(y, 0x7650), b101 = NRC_B2(y, 0x7651), b011 = NRC_B2(y, 0x7652), b111 = NRC_B2(y, 0x7653);
#undef NRC_B2
    const f2pair kk = pk2(-8388608.0f, -8388608.0f);
    const f2pair x00 = fma2(a, sub2(b100, b000), add2(b000, kk)), x10 = fma2(a, sub2(b110, b010), add2(b010, kk));
    const f2pair x01 = fma2(a, sub2(b101, b001), add2(b001, kk)), x11 = fma2(a, sub2(b111, b011), add2(b011, kk));
    const f2pair y0 = fma2(b, sub2(x10, x00), x00), y1 = fma2(b, sub2(x11, x01), x01);
    return fma2(c, sub2(y1, y0), y0);
}

// GetTransmittance(start, end, count) (nrc-forward.frag:641-662): prod exp(-d_i*step) = exp(-step * sum d_i).
// One property of the reference's product IS control flow: the fragment shader replaces the pixel by the
// environment colour when the path transmittance is exactly 1.0 (fwd.frag:859-863), and the product is
// exactly 1.0 iff every factor rounds to 1.0 -- which happens for real (thin fringe of the cloud, short segments).
// So: return exactly 1 when the LARGEST per-tap exponent still rounds to exp() == 1 under nrc_exp, otherwise a
// value strictly below 1.
//
//
// Measured and NOT adopted (the code is gone): taps through the texture unit (7 instead of 32 instructions per tap, but the
// L1/TEX pipe that bounds the light kernels pays twice as much per trilinear fetch) and occupancy-tested taps (a third of the brick loads
// skipped, 8 more instructions per tap on an issue-bound frame).
template <int COUNT>
__device__ __forceinline__ float transmittance(const TapCtx& t, float densityFactor, f3 a, f3 b, unsigned& taps)
{
    float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
    float stepSize = sqrtf(dx * dx + dy * dy + dz * dz) / (float)COUNT;
    if (stepSize == 0.0f) return 1.0f;
    const float u0 = fmaf(a.x, t.sx, t.ox), v0 = fmaf(a.y, t.sy, t.oy), s0 = fmaf(a.z, t.sz, t.oz);
    const float u1 = fmaf(b.x, t.sx, t.ox), v1 = fmaf(b.y, t.sy, t.oy), s1 = fmaf(b.z, t.sz, t.oz);
    const float du = (u1 - u0) * (1.0f / (float)COUNT), dv = (v1 - v0) * (1.0f / (float)COUNT), ds = (s1 - s0) * (1.0f / (float)COUNT);
    const bool inside = (fminf(u0, u1) > -1.0f) & (fmaxf(u0, u1) < t.nx) & (fminf(v0, v1) > -1.0f) & (fmaxf(v0, v1) < t.ny) &
                        (fminf(s0, s1) > -1.0f) & (fmaxf(s0, s1) < t.nz);
    float sum = 0.0f, dmax = 0.0f;
    if (inside) {
        const f2pair du2 = pk2(du, du), dv2 = pk2(dv, dv), ds2 = pk2(ds, ds), u02 = pk2(u0, u0), v02 = pk2(v0, v0), s02 = pk2(s0, s0);
        const f2pair two = pk2(2.0f, 2.0f);
        f2pair fi = pk2(0.0f, 1.0f), sum2 = pk2(0.0f, 0.0f);
#pragma unroll 1
        for (int i = 0; i < COUNT; i += TAP_BATCH) {
            uint2 q[TAP_BATCH];
            f2pair fa[TAP_BATCH / 2], fb[TAP_BATCH / 2], fc[TAP_BATCH / 2];
#pragma unroll
            for (int j = 0; j < TAP_BATCH / 2; j++) {
                int c0, c1;
                tap_cell2(t, fma2(fi, du2, u02), fma2(fi, dv2, v02), fma2(fi, ds2, s02), c0, c1, fa[j], fb[j], fc[j]);
                q[2 * j] = __ldg(t.bricks + c0);
                q[2 * j + 1] = __ldg(t.bricks + c1);
                fi = add2(fi, two);
            }
#pragma unroll
            for (int j = 0; j < TAP_BATCH / 2; j++) {
                const f2pair d2 = tap_blend2(q[2 * j], q[2 * j + 1], fa[j], fb[j], fc[j]);
                sum2 = add2(sum2, d2);
                float d0, d1;
                upk2(d2, d0, d1);
                dmax = fmaxf(dmax, fmaxf(d0, d1));
            }
        }
        float se, so;
        upk2(sum2, se, so);
        sum = se + so;
    } else {   // segment leaves the sampled region (or is not finite): test every tap
        float fi = 0.0f;
#pragma unroll 1
        for (int i = 0; i < COUNT; i++) {
            float u = fmaf(fi, du, u0), v = fmaf(fi, dv, v0), s = fmaf(fi, ds, s0);
            fi += 1.0f;
            bool in = (u > -1.0f) & (u < t.nx) & (v > -1.0f) & (v < t.ny) & (s > -1.0f) & (s < t.nz);
            float d = in ? tap_texel(t, u, v, s) : 0.0f;
            sum += d;
            dmax = fmaxf(dmax, d);
        }
    }
    taps += COUNT;
    const float k = densityFactor * (1.0f / 255.0f);
    float argMax = (dmax * k) * stepSize;
    if (argMax < 1.0e-6f && nrc_exp(-argMax) == 1.0f) return 1.0f;
    return fminf(__expf(-(sum * k) * stepSize), 0.99999994f);
}

// 8 resident blocks are REQUIRED of ptxas: without a minimum it chases occupancy and sinks every brick load next to its use (one
// load in flight per thread, ncu r01f: 35-57 % long-scoreboard); with a 64-register budget it keeps the batch of loads together.
template <bool RENDER, bool CDF>
__global__ void __launch_bounds__(128, 8) k_light(const __grid_constant__ TraceParams P, VertexPool pool, unsigned long long* tapCounter)
{
    const uint32_t count = min(*pool.count, pool.cap);
    const Box box = make_box(P);
    const TapCtx tc = make_tap(P);
    const float kDensity = P.densityFactor;
    unsigned taps = 0;
    for (uint32_t vtx = blockIdx.x * blockDim.x + threadIdx.x; vtx < count; vtx += gridDim.x * blockDim.x) {
        float4 A = pool.a[vtx];
        if (A.w < 0.0f) continue;   // unused slot of a train-walk chunk
        float4 B = pool.b[vtx], C = pool.c[vtx];
        f3 pos = mk3(A.x, A.y, A.z), dir = mk3(B.x, B.y, B.z), last = mk3(C.x, C.y, C.z);
        f3 light = mk3(0.0f, 0.0f, 0.0f);
        unsigned sdf = 0;
        if (P.dirStrength != 0.0f) {   // TraceDirLight (fwd.frag:664-675), colour ignored (A21)
            f3 L = mk3(P.dirDir[0], P.dirDir[1], P.dirDir[2]);
            f3 en, ex;
            find_entry_exit(box, pos, s_neg(s_normalize(L)), en, ex, sdf);
            float T = transmittance<32>(tc, kDensity, pos, ex, taps);
            float l = T * P.dirStrength * hg_phase(P.g, -dot3(L, dir));
            light.x += l; light.y += l; light.z += l;
        }
        if (P.ptStrength != 0.0f) {    // TracePointLight (fwd.frag:677-688)
            f3 lp = mk3(P.ptPos[0], P.ptPos[1], P.ptPos[2]);
            float T = transmittance<32>(tc, kDensity, lp, pos, taps);
            f3 tl = mk3(lp.x - pos.x, lp.y - pos.y, lp.z - pos.z);
            float inv = rsqrtf(dot3(tl, tl));
            float ph = hg_phase(P.g, -dot3(tl, dir) * inv);
            float l = P.ptStrength * T * ph;
            light.x += P.ptColor[0] * l; light.y += P.ptColor[1] * l; light.z += P.ptColor[2] * l;
        }
        if (RENDER) {                  // SampleHdrEnvMap(pos, dir, 8) (fwd.frag:710-749), first loop only
            f3 acc = mk3(0.0f, 0.0f, 0.0f);
#pragma unroll 1
            for (int k = 0; k < 4; k++) {
                float4 r = pool.rnd[(size_t)vtx * 4 + k];
#pragma unroll 1
                for (int h = 0; h < 2; h++) {
                    const float r0 = h ? r.z : r.x, r1 = h ? r.w : r.y;
                    f3 rdir, c;
                    float phase = 1.0f;
                    if (CDF && k >= 2) {   // opt-in CDF half (fwd.frag:727-744): samples 4..7, draws 8..15
                        float thetaNorm = tex1_clamp(P.envInvCdfY, P.envH, r0);
                        float phiNorm = tex2_clamp(P.envInvCdfX, P.envW, P.envH, r1, thetaNorm);
                        // the unit direction the map's uv convention (fwd.frag:690-708) assigns to texel (phiNorm, thetaNorm); the
                        // expression at fwd.frag:735 is neither unit length nor that direction (dead code there, NaN phase if run)
                        float th = thetaNorm * 3.14159265358979323846f, ph2 = phiNorm * 2.0f * 3.14159265358979323846f;
                        float st = sinf(th);
                        rdir = mk3(-st * cosf(ph2), -cosf(th), -st * sinf(ph2));
                        phase = hg_phase(P.g, -dot3(rdir, dir));
                        c = env_texel(P, phiNorm, thetaNorm, P.envHpm);
                    } else {
                        rdir = new_ray_dir_given(dir, P.g, r0, r1);
                        c = env_lookup(P, rdir, P.envHpm);
                    }
                    f3 en, ex;
                    find_entry_exit(box, pos, rdir, en, ex, sdf);
                    float T = transmittance<16>(tc, kDensity, pos, ex, taps) * phase;
                    acc.x += c.x * T; acc.y += c.y * T; acc.z += c.z * T;
                }
            }
            light.x += acc.x * 0.125f; light.y += acc.y * 0.125f; light.z += acc.z * 0.125f;
        }
        float tr = transmittance<32>(tc, kDensity, pos, last, taps);
        pool.out[vtx] = make_float4(light.x, light.y, light.z, tr);
    }
    if (tapCounter) {
        unsigned long long t = taps;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(tapCounter, t);
    }
}

// ---- combine kernels: fold vertices in path order -------------------------------------------------------------------
__device__ __forceinline__ void fold_path(const VertexPool& pool, int head, f3& L, float& T)
{
    L = mk3(0.0f, 0.0f, 0.0f);
    T = 1.0f;
    int v = head;
    while (v >= 0) {
        float4 o = pool.out[v];
        float dens = pool.a[v].w;
        L.x += T * (dens * o.x); L.y += T * (dens * o.y); L.z += T * (dens * o.z);   // scatteredLight += T * (density * light)
        T *= o.w;                                                                    // transmittance *= t_r
        v = __float_as_int(pool.b[v].w);
    }
}

__global__ void __launch_bounds__(128) k_render_combine(uint32_t width, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                                                        VertexPool pool, RenderPaths paths, QueryList q)
{
    uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= rows * width * spp) return;
    uint32_t rowLen = width * spp;
    size_t path = (size_t)(y0 + (local / rowLen) * ystride) * rowLen + (local % rowLen);
    f3 L; float T;
    fold_path(pool, paths.head[path], L, T);
    paths.result[path] = make_float4(L.x, L.y, L.z, T);
    float4 qd = paths.qdirValid[path];
    int slot = -1;
    float w = 0.0f;
    if (qd.w != 0.0f) {
        float4 qp = paths.qposPhase[path];
        w = T * qp.w;
        uint32_t s = alloc_slot(q.count);
        if (s < q.cap) {
            slot = (int)s;
            q.pos[3 * s] = qp.x; q.pos[3 * s + 1] = qp.y; q.pos[3 * s + 2] = qp.z;
            q.dir[3 * s] = qd.x; q.dir[3 * s + 1] = qd.y; q.dir[3 * s + 2] = qd.z;
        }
    }
    paths.qweight[path] = w;
    paths.qslot[path] = slot;
}

// One thread per (sample, path): the 8 folds of a sample run in 8 adjacent lanes (each is a pointer chase through the
// vertex chain, latency-bound), then lane 0 of the group adds them up in path order like the reference's loop
// (nrc-train.comp:1217-1223).
__global__ void __launch_bounds__(128) k_train_combine(uint32_t nLocal, VertexPool pool, TrainPaths paths,
                                                       const unsigned long long* __restrict__ taps, unsigned long long* __restrict__ totals)
{
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i = t >> 3, p = t & 7u;
    if (t == 0 && totals) {   // running totals of this rank (nrc_frame_totals)
        totals[1] += nLocal;
        totals[3] += taps[0];
        totals[5] += min(*pool.count, pool.cap);
    }
    f3 L = mk3(0.0f, 0.0f, 0.0f);
    float T = 1.0f;
    if (i < nLocal) fold_path(pool, paths.head[(size_t)i * 8 + p], L, T);
    f3 sum = mk3(0.0f, 0.0f, 0.0f);
    const unsigned base = (threadIdx.x & 31u) & ~7u;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        sum.x += __shfl_sync(0xffffffffu, L.x, base + k);
        sum.y += __shfl_sync(0xffffffffu, L.y, base + k);
        sum.z += __shfl_sync(0xffffffffu, L.z, base + k);
    }
    if (i < nLocal && p == 0) {
        paths.target[3 * (size_t)i] = sum.x / 8.0f;
        paths.target[3 * (size_t)i + 1] = sum.y / 8.0f;
        paths.target[3 * (size_t)i + 2] = sum.z / 8.0f;
    }
}

// fwd.frag:824-870: average the spp paths, add the NRC radiance, env-map colour for misses and T == 1 (A20).
__global__ void __launch_bounds__(128) k_composite(uint32_t width, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                                                   RenderPaths paths, QueryList q,
                                                   const float4* __restrict__ background, float4* __restrict__ frame,
                                                   uint32_t* __restrict__ frame8, float* __restrict__ queryProbe,
                                                   const uint32_t* __restrict__ scalars, const unsigned long long* __restrict__ taps,
                                                   unsigned long long* __restrict__ totals)
{
    uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= rows * width) return;
    if (local == 0 && totals) {   // running totals of this rank (nrc_frame_totals)
        totals[0] += min(*q.count, q.cap);
        totals[2] += taps[0];
        totals[4] += scalars[0];
    }
    size_t pix = (size_t)(y0 + (local / width) * ystride) * width + (local % width);
    float4 bg = background[pix];
    float4 outc;
    if (bg.w != 0.0f) {
        outc = make_float4(bg.x, bg.y, bg.z, 1.0f);
    } else {
        float4 avg = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (uint32_t s = 0; s < spp; s++) {
            size_t path = pix * spp + s;
            float4 r = paths.result[path];
            int slot = paths.qslot[path];
            if (slot >= 0) {
                float w = paths.qweight[path];
                r.x += w * q.rgb[3 * slot]; r.y += w * q.rgb[3 * slot + 1]; r.z += w * q.rgb[3 * slot + 2];
            }
            avg.x += r.x; avg.y += r.y; avg.z += r.z; avg.w += r.w;
        }
        float inv = (float)spp;
        avg.x /= inv; avg.y /= inv; avg.z /= inv; avg.w /= inv;
        outc = (avg.w == 1.0f) ? make_float4(bg.x, bg.y, bg.z, 1.0f) : avg;
    }
    frame[pix] = outc;
    if (frame8) {   // B8G8R8A8_UNORM store: clamp to [0,1], round to nearest
        uint32_t r = (uint32_t)__float2int_rn(__saturatef(outc.x) * 255.0f), g = (uint32_t)__float2int_rn(__saturatef(outc.y) * 255.0f);
        uint32_t b = (uint32_t)__float2int_rn(__saturatef(outc.z) * 255.0f), a = (uint32_t)__float2int_rn(__saturatef(outc.w) * 255.0f);
        frame8[pix] = b | (g << 8) | (r << 16) | (a << 24);
    }
    if (queryProbe) {   // last spp path's query, like the oracle's lastQuery
        size_t path = pix * spp + (spp - 1);
        float4 qp = paths.qposPhase[path], qd = paths.qdirValid[path];
        float* o = queryProbe + 8 * pix;
        o[0] = qp.x; o[1] = qp.y; o[2] = qp.z; o[3] = qd.x; o[4] = qd.y; o[5] = qd.z;
        o[6] = paths.qweight[path]; o[7] = qd.w;
    }
}

// ---- volume: UNORM8 quantisation + 2x2x2 corner bricks (src/Texture3D.cpp:25-40) -----------------------------------
__device__ __forceinline__ uint32_t quant_texel(const float* __restrict__ d, int i, int j, int k, int nx, int ny, int nz)
{
    if (i < 0 || j < 0 || k < 0 || i >= nx || j >= ny || k >= nz) return 0u;   // CLAMP_TO_BORDER, opaque black
    float v = __fmul_rn(d[((size_t)i * ny + j) * nz + k], 255.0f);             // file order x*ny*nz + y*nz + z
    return (uint32_t)((int)v) & 0xffu;                                          // static_cast<uint8_t>(value * 255.0f)
}

__global__ void __launch_bounds__(256) k_build_bricks(const float* __restrict__ d, int nx, int ny, int nz, uint2* __restrict__ bricks)
{
    size_t cells = (size_t)(nx + 1) * (ny + 1) * (nz + 1);
    size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cells) return;
    int ci = (int)(c % (nx + 1)), cj = (int)((c / (nx + 1)) % (ny + 1)), ck = (int)(c / ((size_t)(nx + 1) * (ny + 1)));
    int i0 = ci - 1, j0 = cj - 1, k0 = ck - 1;
    uint32_t lo = 0, hi = 0;
    for (int t = 0; t < 4; t++) {
        lo |= quant_texel(d, i0 + (t & 1), j0 + (t >> 1), k0, nx, ny, nz) << (8 * t);
        hi |= quant_texel(d, i0 + (t & 1), j0 + (t >> 1), k0 + 1, nx, ny, nz) << (8 * t);
    }
    bricks[c] = make_uint2(lo, hi);
}

__global__ void __launch_bounds__(256) k_volume_sample(Grid g, const float* __restrict__ pos, size_t n, float* __restrict__ out)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = grid_density(g, mk3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]));
}

// ---- launchers -----------------------------------------------------------------------------------------------------------
nrc_status launch_build_bricks(nrc_volume* vol, const float* densityDev)
{
    size_t cells = (size_t)(vol->nx + 1) * (vol->ny + 1) * (vol->nz + 1);
    cudaStream_t st = vol->ctx->stream;
    NRC_LAUNCH(vol->ctx, st, K_BUILD_BRICKS,
               k_build_bricks<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(densityDev, (int)vol->nx, (int)vol->ny, (int)vol->nz, vol->bricks));
    return NRC_OK;
}

nrc_status launch_volume_sample(nrc_volume* vol, const float* pos, size_t n, float* out)
{
    if (n == 0) return NRC_OK;
    Grid g;
    g.nx = (int)vol->nx; g.ny = (int)vol->ny; g.nz = (int)vol->nz;
    g.bricks = vol->bricks;
    g.skySize = mk3(62.5f, 42.5f, 76.5f);
    g.invSize = mk3(1.0f / 62.5f, 1.0f / 42.5f, 1.0f / 76.5f);   // IEEE single division on the host = __frcp_rn
    g.densityFactor = vol->params.densityFactor;
    cudaStream_t st = vol->ctx->stream;
    NRC_LAUNCH(vol->ctx, st, K_VOLUME_SAMPLE, k_volume_sample<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g, pos, n, out));
    return NRC_OK;
}

// The tracing kernels use no shared memory, the MLP kernels 61-224 KB per CTA.  Left to the driver's default, an SM that has run an MLP
// CTA keeps its large shared-memory split for the tracing CTAs that follow, which then gather through a fraction of the L1 (forcing the
// largest split on them costs 40 %: 4.19 against 2.91 ms per 1080p frame).  An explicit preference hands them the L1 back: any value from
// 5 to 28 % measures the same (1080p 2.91 -> 2.81 ms, 720p 1.97 -> 1.80 ms, profiles/r02_carveout.md); 25 % = 64 KB still takes one
// CTA of the inference kernel beside them.  NRC_CARVEOUT=pct overrides, -1 leaves the driver's default.  Once per device.
static void trace_carveout(nrc_ctx* ctx)
{
    static bool done[64] = {};
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64 || done[dev]) return;
    done[dev] = true;
    const char* e = getenv("NRC_CARVEOUT");
    const int pct = e ? atoi(e) : 25;
    if (pct < 0 || pct > 100) return;
#define CARVE(k) cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, pct)
    CARVE(k_render_walk<8>); CARVE(k_render_walk_tail<8>);
    CARVE(k_train_walk<true>); CARVE(k_train_walk<false>);
    CARVE((k_light<true, true>)); CARVE((k_light<true, false>)); CARVE((k_light<false, false>));
    CARVE(k_render_combine); CARVE(k_train_combine); CARVE(k_composite);
#undef CARVE
    cudaGetLastError();   // a hint: nothing to report if the driver declines it
}

nrc_status launch_render_walk(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                              VertexPool pool, RenderPaths paths, float4* background, uint32_t* counters)
{
    uint32_t n = rows * P.width;
    if (n == 0) return NRC_OK;
    trace_carveout(ctx);
    // steps in the head launch; 0 = the whole walk in one launch.  Read at every launch (tests switch it)
    const char* se = getenv("NRC_RWALK_SPLIT");
    int splitEnv = se ? atoi(se) : 10;
    if (splitEnv < 0 || splitEnv >= NRC_RENDER_WALK_STEPS) splitEnv = 0;
    // the caller zeroes *paths.spillCount on this stream before the call (renderer.cu: render_pass)
    const int split = (spp == 1 && paths.spill != nullptr && paths.spillStride >= n) ? splitEnv : 0;
    const unsigned blocks = (n + 127) / 128;   // the tail launch: one thread per spilled path
    const unsigned headBlocks = (((P.width + 7u) / 8u) * ((rows + 3u) / 4u) * 32u + 127u) / 128u;   // the head launch: 8 x 4 pixel tiles
#define RWALK_CASE(MBV)                                                                                                                    \
    NRC_LAUNCH(ctx, st, K_RENDER_WALK,                                                                                                     \
               (k_render_walk<MBV><<<headBlocks, 128, 0, st>>>(P, y0, rows, ystride, spp, pool, paths, background, counters, split))); \
    if (split > 0)                                                                                                                         \
        NRC_LAUNCH(ctx, st, K_RENDER_WALK_TAIL, (k_render_walk_tail<MBV><<<blocks, 128, 0, st>>>(P, pool, paths, counters, split)));
    RWALK_CASE(8)   // 64 registers: 8 resident blocks, the fastest of the budgets tried (profiles/r01j)
#undef RWALK_CASE
    return NRC_OK;
}

nrc_status launch_train_walk(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t i0, uint32_t n, uint32_t istride, VertexPool pool,
                             TrainPaths paths, uint32_t* counters)
{
    if (n == 0) return NRC_OK;
    trace_carveout(ctx);
    const uint32_t totalWarps = (n + 31u) / 32u;
    const uint32_t warpsPerBlock = 4;
    const uint32_t blocks = (totalWarps + warpsPerBlock - 1) / warpsPerBlock;
    const size_t smem = 0;
    if (pool.rnd != nullptr)   // the pool carries random numbers: env-map term in the training targets is on (renderer.cu)
        NRC_LAUNCH(ctx, st, K_TRAIN_WALK,
                   k_train_walk<true><<<blocks, 32 * warpsPerBlock, smem, st>>>(P, i0, n, istride, pool, paths, counters));
    else
        NRC_LAUNCH(ctx, st, K_TRAIN_WALK,
                   k_train_walk<false><<<blocks, 32 * warpsPerBlock, smem, st>>>(P, i0, n, istride, pool, paths, counters));
    return NRC_OK;
}

// render = with the env-map gather (needs pool.rnd); kid = which profile slot the launch is booked under
nrc_status launch_light(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, VertexPool pool, bool render, unsigned long long* tapCounter, int kid)
{
    // the vertex count lives on the device: a grid-stride launch sized for the machine, not for the count
    unsigned want = (pool.cap + 127) / 128;
    // 64 blocks per SM (8 are resident at a time): the lights run beside the walks of other frames, and many short-lived blocks share the
    // SMs with them more evenly than two waves of long-lived ones (16 per SM: 2.656 ms per 1080p frame; 32: 2.62; 64: 2.604; 128: 2.607;
    // 8 and below lose).  NRC_LIGHT_BLOCKS_PER_SM=n overrides.
    static int perSm = 0;
    if (perSm == 0) { const char* le = getenv("NRC_LIGHT_BLOCKS_PER_SM"); perSm = le ? atoi(le) : 64; if (perSm < 1) perSm = 64; }
    unsigned cap = (unsigned)ctx->smCount * (unsigned)perSm;
    unsigned blocks = want < cap ? want : cap;
    if (blocks == 0) return NRC_OK;
    if (render && P.envInvCdfX) NRC_LAUNCH(ctx, st, kid, (k_light<true, true><<<blocks, 128, 0, st>>>(P, pool, tapCounter)));
    else if (render) NRC_LAUNCH(ctx, st, kid, (k_light<true, false><<<blocks, 128, 0, st>>>(P, pool, tapCounter)));
    else NRC_LAUNCH(ctx, st, kid, (k_light<false, false><<<blocks, 128, 0, st>>>(P, pool, tapCounter)));
    return NRC_OK;
}

nrc_status launch_render_combine(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                                 VertexPool pool, RenderPaths paths, QueryList q)
{
    uint32_t n = rows * P.width * spp;
    if (n == 0) return NRC_OK;
    NRC_LAUNCH(ctx, st, K_RENDER_COMBINE,
               k_render_combine<<<(n + 127) / 128, 128, 0, st>>>(P.width, y0, rows, ystride, spp, pool, paths, q));
    return NRC_OK;
}

nrc_status launch_train_combine(nrc_ctx* ctx, cudaStream_t st, uint32_t nLocal, VertexPool pool, TrainPaths paths,
                                const unsigned long long* taps, unsigned long long* totals)
{
    if (nLocal == 0) return NRC_OK;
    NRC_LAUNCH(ctx, st, K_TRAIN_COMBINE, k_train_combine<<<(nLocal * 8 + 127) / 128, 128, 0, st>>>(nLocal, pool, paths, taps, totals));
    return NRC_OK;
}

nrc_status launch_composite(nrc_ctx* ctx, cudaStream_t st, const TraceParams& P, uint32_t y0, uint32_t rows, uint32_t ystride, uint32_t spp,
                            RenderPaths paths,
                            QueryList q, const float4* background, float4* frame, uint32_t* frame8, float* queryProbe,
                            const uint32_t* scalars, const unsigned long long* taps, unsigned long long* totals)
{
    uint32_t n = rows * P.width;
    if (n == 0) return NRC_OK;
    NRC_LAUNCH(ctx, st, K_COMPOSITE,
               k_composite<<<(n + 127) / 128, 128, 0, st>>>(P.width, y0, rows, ystride, spp, paths, q, background, frame, frame8, queryProbe, scalars,
                                                            taps, totals));
    return NRC_OK;
}
