// walk_host.cpp -- TEST HARNESS: compiles neuralradiancecaching_b200/csrc/{strict,walk}.cuh with g++ so the device
// walk code can be compared with the oracle on the (GPU-less) build machine.  Not linked into the product.
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../neuralradiancecaching_b200/csrc/walk.cuh"

struct RecEmitter {
    std::vector<float>* out;   // per vertex: pos3, density, dir3, last3
    unsigned n;
    bool vertex(f3 pos, float density, f3 dir, f3 last, const float*) {
        if (out) { float v[10] = {pos.x, pos.y, pos.z, density, dir.x, dir.y, dir.z, last.x, last.y, last.z}; out->insert(out->end(), v, v + 10); }
        n++;
        return true;
    }
};

static std::vector<nrc_brick_t> g_bricks;

extern "C" {

// density in file order x*ny*nz + y*nz + z; same quantisation/brick layout as k_build_bricks
void emu_set_volume(const float* d, int nx, int ny, int nz)
{
    g_bricks.assign((size_t)(nx + 1) * (ny + 1) * (nz + 1), nrc_brick_t{0, 0});
    auto tex = [&](int i, int j, int k) -> uint32_t {
        if (i < 0 || j < 0 || k < 0 || i >= nx || j >= ny || k >= nz) return 0u;
        float v = d[((size_t)i * ny + j) * nz + k] * 255.0f;
        return (uint32_t)((int)v) & 0xffu;
    };
    for (int ck = 0; ck <= nz; ck++) for (int cj = 0; cj <= ny; cj++) for (int ci = 0; ci <= nx; ci++) {
        uint32_t lo = 0, hi = 0;
        for (int t = 0; t < 4; t++) {
            lo |= tex(ci - 1 + (t & 1), cj - 1 + (t >> 1), ck - 1) << (8 * t);
            hi |= tex(ci - 1 + (t & 1), cj - 1 + (t >> 1), ck) << (8 * t);
        }
        g_bricks[((size_t)ck * (ny + 1) + cj) * (nx + 1) + ci] = nrc_brick_t{lo, hi};
    }
}

static void setup(const float* invProjView, const float* camPos, const float* random2, float g, unsigned useNN, unsigned showNonNN,
                  unsigned W, unsigned H, unsigned trainW, unsigned trainH, float densityFactor, int nx, int ny, int nz,
                  WalkParams& P, Box& box, Grid& grid)
{
    memcpy(P.invProjView, invProjView, 64); memcpy(P.camPos, camPos, 12);
    P.randomX = random2[0]; P.randomY = random2[1]; P.g = g; P.useNN = useNN; P.showNonNN = showNonNN;
    P.width = W; P.height = H; P.trainW = trainW;
    P.widthFactor = 1.0f / (float)trainW; P.heightFactor = 1.0f / (float)trainH;
    box.half = mk3(31.25f, 21.25f, 38.25f);
    box.exitJump = __builtin_sqrtf(46259.0f);
    grid.nx = nx; grid.ny = ny; grid.nz = nz; grid.bricks = g_bricks.data();
    grid.skySize = mk3(62.5f, 42.5f, 76.5f); grid.densityFactor = densityFactor;
}

// counters[H*W][4] = draws, lit, steps, emitted ; query[H*W][7] = pos, dir, valid
void emu_render_walk(const float* invProjView, const float* camPos, const float* random2, float g, unsigned useNN, unsigned showNonNN,
                     unsigned W, unsigned H, unsigned spp, float densityFactor, int nx, int ny, int nz, unsigned* counters, float* query)
{
    WalkParams P; Box box; Grid grid;
    setup(invProjView, camPos, random2, g, useNN, showNonNN, W, H, 1, 1, densityFactor, nx, ny, nz, P, box, grid);
#pragma omp parallel for schedule(dynamic, 1)
    for (long py = 0; py < (long)H; py++) for (unsigned px = 0; px < W; px++) {
        f3 ro, rd; Rng rng;
        render_ray(P, px, (unsigned)py, ro, rd, rng);
        WalkCounters wc = {0, 0, 0, 0};
        f3 entry, ex;
        find_entry_exit(box, ro, rd, entry, ex, wc.sdf);
        bool miss = sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE;
        WalkQuery q; q.valid = false; q.pos = q.dir = q.lastDir = mk3(0, 0, 0);
        RecEmitter em{nullptr, 0};
        if (!miss) for (unsigned s = 0; s < spp; s++) render_walk_path(P, box, grid, entry, rd, rng, em, wc, q);
        size_t p = (size_t)py * W + px;
        counters[4 * p] = rng.draws; counters[4 * p + 1] = wc.lit; counters[4 * p + 2] = wc.steps; counters[4 * p + 3] = wc.emitted;
        float* o = query + 7 * p;
        o[0] = q.pos.x; o[1] = q.pos.y; o[2] = q.pos.z; o[3] = q.dir.x; o[4] = q.dir.y; o[5] = q.dir.z; o[6] = q.valid ? 1.0f : 0.0f;
    }
}

// The render walk as the two launches of trace.cu run it: `splitAt` steps, the state of a surviving path packed into the 20-word spill
// record (same fields, same order as k_render_walk writes them), unpacked, the remaining steps.  Outputs as emu_render_walk (spp = 1).
void emu_render_walk_split(const float* invProjView, const float* camPos, const float* random2, float g, unsigned useNN, unsigned showNonNN,
                           unsigned W, unsigned H, int splitAt, float densityFactor, int nx, int ny, int nz, unsigned* counters, float* query,
                           unsigned* spilled)
{
    WalkParams P; Box box; Grid grid;
    setup(invProjView, camPos, random2, g, useNN, showNonNN, W, H, 1, 1, densityFactor, nx, ny, nz, P, box, grid);
    unsigned nSpilled = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : nSpilled)
    for (long py = 0; py < (long)H; py++) for (unsigned px = 0; px < W; px++) {
        f3 ro, rd; Rng rng;
        render_ray(P, px, (unsigned)py, ro, rd, rng);
        WalkCounters wc = {0, 0, 0, 0};
        f3 entry, ex;
        find_entry_exit(box, ro, rd, entry, ex, wc.sdf);
        bool miss = sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE;
        WalkQuery q; q.valid = false; q.pos = q.dir = q.lastDir = mk3(0, 0, 0);
        RecEmitter em{nullptr, 0};
        if (!miss) {
            WalkState ws = walk_state_begin(entry, rd);
            if (render_walk_steps(P, box, grid, ws, 0, splitAt, rng, em, wc, q)) {
                float rec[20] = {ws.cur.x, ws.cur.y, ws.cur.z, ws.termProb, ws.last.x, ws.last.y, ws.last.z, rng.pre,
                                 ws.dir.x, ws.dir.y, ws.dir.z, rng.prePre, ws.lastDir.x, ws.lastDir.y, ws.lastDir.z, 0.0f,
                                 0.0f, 0.0f, 0.0f, 0.0f};
                unsigned draws = rng.draws, lit = wc.lit;
                nSpilled++;
                // ---- tail launch: nothing but the record, the draw count and the lit count survives
                WalkState t;
                t.cur = mk3(rec[0], rec[1], rec[2]); t.termProb = rec[3];
                t.last = mk3(rec[4], rec[5], rec[6]);
                t.dir = mk3(rec[8], rec[9], rec[10]);
                t.lastDir = mk3(rec[12], rec[13], rec[14]);
                Rng r2; r2.pre = rec[7]; r2.prePre = rec[11]; r2.draws = draws;
                WalkCounters w2 = {lit, (unsigned)splitAt, 0, 0};
                WalkQuery q2; q2.valid = false; q2.pos = q2.dir = q2.lastDir = mk3(0, 0, 0);
                render_walk_steps(P, box, grid, t, splitAt, NRC_RENDER_WALK_STEPS, r2, em, w2, q2);
                rng = r2; wc = w2; q = q2;
            }
        }
        size_t p = (size_t)py * W + px;
        counters[4 * p] = rng.draws; counters[4 * p + 1] = wc.lit; counters[4 * p + 2] = wc.steps; counters[4 * p + 3] = wc.emitted;
        float* o = query + 7 * p;
        o[0] = q.pos.x; o[1] = q.pos.y; o[2] = q.pos.z; o[3] = q.dir.x; o[4] = q.dir.y; o[5] = q.dir.z; o[6] = q.valid ? 1.0f : 0.0f;
    }
    *spilled = nSpilled;
}

// Divergence study (scripts/walk_divergence.py): per pixel {walk steps, SDF evaluations of the walk, SDF evaluations of the camera ray}
void emu_render_walk_cost(const float* invProjView, const float* camPos, const float* random2, float g, unsigned useNN, unsigned W, unsigned H,
                          float densityFactor, int nx, int ny, int nz, unsigned* cost, unsigned char* perStep)
{
    WalkParams P; Box box; Grid grid;
    setup(invProjView, camPos, random2, g, useNN, 0, W, H, 1, 1, densityFactor, nx, ny, nz, P, box, grid);
#pragma omp parallel for schedule(dynamic, 1)
    for (long py = 0; py < (long)H; py++) for (unsigned px = 0; px < W; px++) {
        f3 ro, rd; Rng rng;
        render_ray(P, px, (unsigned)py, ro, rd, rng);
        WalkCounters wc = {0, 0, 0, 0};
        f3 entry, ex;
        find_entry_exit(box, ro, rd, entry, ex, wc.sdf);
        unsigned cam = wc.sdf;
        bool miss = sky_sdf(box, entry) > NRC_MAX_RAY_DISTANCE;
        WalkQuery q; q.valid = false;
        RecEmitter em{nullptr, 0};
        size_t p = (size_t)py * W + px;
        if (!miss && perStep) {
            // the same walk one step at a time: SDF evaluations of step i (render_walk_path's loop body, lighting left out)
            const bool nn = (P.useNN == 1u);
            f3 cur = entry, dir = rd;
            float termProb = 1.0f;
            for (int i = 0; i < 32; i++) {
                wc.steps++;
                float density = grid_density(grid, cur);
                if (density > 0.0f) {
                    if (nn) {
                        if (rng_next(rng) > termProb) break;
                        termProb = NRC_MUL(termProb, 0.5f);
                    }
                    for (int k = 0; k < 16; k++) rng_next(rng);
                    wc.lit++;
                    dir = new_ray_dir(dir, P.g, rng);
                }
                unsigned before = wc.sdf;
                cur = walk_advance(box, cur, dir, rng, wc.sdf);
                perStep[32 * p + i] = (unsigned char)(wc.sdf - before > 255 ? 255 : wc.sdf - before);
            }
        } else if (!miss) render_walk_path(P, box, grid, entry, rd, rng, em, wc, q);
        cost[4 * p] = wc.steps; cost[4 * p + 1] = wc.sdf - cam; cost[4 * p + 2] = cam; cost[4 * p + 3] = wc.lit;
    }
}

// records[n][6] = pos, dir ; counters[n][3] = draws, lit, steps
void emu_train_walk(const float* invProjView, const float* camPos, const float* random2, float g, unsigned trainW, unsigned trainH,
                    float densityFactor, int nx, int ny, int nz, float* records, unsigned* counters, int envLight)
{
    WalkParams P; Box box; Grid grid;
    setup(invProjView, camPos, random2, g, 0, 0, 1, 1, trainW, trainH, densityFactor, nx, ny, nz, P, box, grid);
    long n = (long)trainW * trainH;
#pragma omp parallel for schedule(dynamic, 4)
    for (long i = 0; i < n; i++) {
        Rng rng; WalkCounters wc = {0, 0, 0, 0};
        f3 cur, dir;
        train_pick_vertex(P, box, grid, (unsigned)i, rng, wc, cur, dir);
        RecEmitter em{nullptr, 0};
        for (int p = 0; p < 8; p++) {
            if (envLight) train_walk_path_env(P, box, grid, cur, dir, rng, em, wc);
            else train_walk_path(P, box, grid, cur, dir, rng, em, wc);
        }
        float* o = records + 6 * i;
        o[0] = cur.x; o[1] = cur.y; o[2] = cur.z; o[3] = dir.x; o[4] = dir.y; o[5] = dir.z;
        counters[3 * i] = rng.draws; counters[3 * i + 1] = wc.lit; counters[3 * i + 2] = wc.steps;
    }
}

// MRHE encode through the device code path: feat[n][32], idx[n][128]
void emu_encode(const float* pos, long n, const float* tables, unsigned L, unsigned T, const unsigned* res, float* feat, unsigned* idx)
{
    for (long q = 0; q < n; q++) {
        f3 p = mk3(pos[3 * q], pos[3 * q + 1], pos[3 * q + 2]);
        f3 np = mrhe_norm_pos(p, mk3(62.5f, 42.5f, 76.5f));
        for (unsigned l = 0; l < 16; l++) {
            unsigned id[8] = {0}; f3 lerp; float f0[8], f1[8]; float o0 = 0, o1 = 0;
            if (l < L) {
                mrhe_level_indices(p, np, res[l], T, id, lerp);
                for (int c = 0; c < 8; c++) { f0[c] = tables[((size_t)l * T + id[c]) * 2]; f1[c] = tables[((size_t)l * T + id[c]) * 2 + 1]; }
                o0 = mrhe_blend(f0, lerp); o1 = mrhe_blend(f1, lerp);
            }
            feat[32 * q + 2 * l] = o0; feat[32 * q + 2 * l + 1] = o1;
            for (int c = 0; c < 8; c++) idx[128 * q + 8 * l + c] = id[c];
        }
    }
}

void emu_oneblob(const float* dir, long n, float* feat)
{
    for (long q = 0; q < n; q++) oneblob_encode(mk3(dir[3 * q], dir[3 * q + 1], dir[3 * q + 2]), feat + 32 * q, 1);
}

float emu_density(const float* p, float densityFactor, int nx, int ny, int nz)
{
    Grid grid; grid.nx = nx; grid.ny = ny; grid.nz = nz; grid.bricks = g_bricks.data();
    grid.skySize = mk3(62.5f, 42.5f, 76.5f); grid.densityFactor = densityFactor;
    return grid_density(grid, mk3(p[0], p[1], p[2]));
}

}  // extern "C"
