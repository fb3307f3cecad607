// lensrt — the render megakernel: sampler -> sensor (realistic lens / pinhole) -> path integrator -> film splat.
// Replaces SamplingIntegrator::render + render_sample, JIT branch (src/render/integrator.cpp:247-330,419-491),
// PathIntegrator::sample (src/integrators/path.cpp:95-287) and ImageBlock::put + put_block
// (src/render/imageblock.cpp:135-172,174-532) for hdrfilm's R,G,B,W storage (src/films/hdrfilm.cpp:245-297).
//
// Execution model (not the reference's wavefront): persistent CTAs pull work items (pixel tile x sample chunk) from
// a global counter. The unit of work is a WARP: a tile is one row of 32 pixels, the warp owns the pool of its
// 32 x chunk_spp samples and a private shared-memory film tile; the warps of a CTA share only the staged lens table. Sample `j` of pixel `p` always uses PCG32 stream idx = p*spp + j (sampler.cpp:121-130)
// whatever the sharding, so images are identical up to float summation order for any tile / chunk / GPU split.
// The three stages have very different shapes, so each runs at its own width and they meet in shared-memory queues:
//   * camera rays are generated 32 at a time with all lanes active (the lens trace is the same instruction sequence for
//     every lane) and the transmitted ones are pushed into a per-warp FIFO (ray, weight, film position, RNG state);
//   * a lane whose path ended pops the next camera ray, so lanes never wait for each other's paths;
//   * BVH traversal is resumable (lrt_trace.cuh): path segments and shadow rays of different lanes share ONE traversal
//     site, and a burst ends as soon as a quarter of its lanes have finished, so they can shade and fetch new rays;
//   * finished samples go to a per-warp splat queue that is drained 32 entries at a time into the shared-memory film
//     tile (float atomics in shared memory are CAS loops on this architecture: run them at full width, two channels
//     per 64-bit CAS), and the tile is flushed once per work item with global float atomics.
#include <cstddef>
#include <mutex>
#include "lrt_internal.h"
#include "lrt_rng.cuh"
#include "lrt_shade.cuh"
#include "lrt_scene.h"

namespace lrt {

constexpr int kTileW = 32, kTileH = 1, kCtaThreads = 128, kWarps = kCtaThreads / 32;
constexpr int kFifo = 64, kFifoWords = 14;      // camera-ray FIFO per warp: o[3] d[3] maxt weight sx sy rng.state rng.inc
constexpr int kQueue = 96, kQueueWords = 5;     // splat queue per warp: sx sy r g b. Drained (32 at a time) at ONE place, the top
                                                // of the warp's loop, where it holds < 32 entries afterwards; camera-ray
                                                // generation and shading add at most 32 each before the next drain
constexpr int kMaxRec = 16;                     // adjoint: bounces of one path that can receive a BSDF gradient
constexpr int kGradSlots = 1024;                // adjoint: per-CTA shared accumulators (6 per BSDF: columns 2..4 and 11..13 of its row; 3 per emitter)

struct RenderArgs {
    SceneView sc;
    const LensSystem *lens;
    FilmParams film;
    uint32_t seed; int spp, max_depth, rr_depth, hide_emitters;
    // base work items = (pixel tile, sample chunk), tile-major. The LAST `n_split` base items are handed out in `n_sub`
    // pieces each (guided self-scheduling): when the work counter runs out, the items still in flight are short, so the
    // tail of the launch is a fraction of a tile's work while all but a few per cent of the samples run in full-size items
    int tiles_x, n_chunks, chunk_spp, n_sub;
    int shard_index, shard_count; long long n_items, n_whole;
    int pad, smem_w, smem_h, exit_drop, leaf_min;
    unsigned long long *counter;
    float *accum;
    int lens_bytes;                // bytes of the LensSystem staged into shared memory (header + interfaces in use)
    // adjoint (k_render<true>): d(loss)/d(storage) in, parameter gradients out (rows laid out like the scene tables)
    const float *grad_storage; float *grad_bsdfs, *grad_emitters; int n_bsdfs;
    unsigned long long *stats;     // [0] samples, [1] lens misses, [2] rays, [3] bounces, [4] BVH nodes fetched, [5] triangles tested, [6] traversal-stack clamps
};

// Shared-memory float atomics are CAS loops on sm_100 (ATOMS.CAST.SPIN): add two neighbouring channels with ONE 64-bit CAS.
LRT_D void atomic_add2_shared(float *p, float a, float b) {
    unsigned long long *q = reinterpret_cast<unsigned long long *>(p);
    unsigned long long old = *q, assumed;
    do {
        assumed = old;
        const float x = __uint_as_float((unsigned) assumed) + a, y = __uint_as_float((unsigned) (assumed >> 32)) + b;
        old = atomicCAS(q, assumed, (unsigned long long) __float_as_uint(x) | ((unsigned long long) __float_as_uint(y) << 32));
    } while (old != assumed);
}

// ImageBlock::put of one sample into the shared-memory tile (tensor coordinates of tile(0,0) = (ox, oy))
LRT_D void splat_to_tile(const FilmParams &F, float *tile, int smem_w, int ox, int oy, float sx, float sy, float3 value) {
    Footprint fp;
    if (!film_footprint(F, sx, sy, fp)) return;
    const int sw = F.size_x + 2 * F.border, shh = F.size_y + 2 * F.border;   // (box filters have no border)
    for (int iy = 0; iy < fp.ny; ++iy) {
        const int y = fp.y0 + iy;
        if ((unsigned) y >= (unsigned) shh) continue;
        for (int ix = 0; ix < fp.nx; ++ix) {
            const int x = fp.x0 + ix;
            if ((unsigned) x >= (unsigned) sw) continue;
            const float w = fp.wy[iy] * fp.wx[ix];
            float *dst = tile + ((y - oy) * smem_w + (x - ox)) * 4;
            if (value.x != 0.f || value.y != 0.f) atomic_add2_shared(dst, value.x * w, value.y * w);
            atomic_add2_shared(dst + 2, value.z * w, w);
        }
    }
}

#ifndef LRT_RENDER_MIN_BLOCKS
#define LRT_RENDER_MIN_BLOCKS 4
#endif
// kLevel 1 (kAttr): the scene has vertex normals / texture coordinates / textures / mesh emitters (lrt_shade.cuh); kLevel 2
// (kAnalytic) adds analytic spheres and the environment emitter. Without them that code is compiled out so that the hot loop
// keeps its instruction-cache footprint and its registers (level 0 is the headline variant).
template <bool kAdj, int kLevel>
__global__ void __launch_bounds__(kCtaThreads, LRT_RENDER_MIN_BLOCKS) k_render(RenderArgs A) {
    constexpr bool kAttr = kLevel >= 1, kAnalytic = kLevel >= 2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LensSystem &L = *reinterpret_cast<LensSystem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int tile_elems = A.smem_w * A.smem_h * 4;
    // per-warp private storage: film tile, camera-ray FIFO [word][slot], splat queue [word][slot]
    float *tile = reinterpret_cast<float *>(smem_raw + A.lens_bytes) + warp * (tile_elems + kFifo * kFifoWords + kQueue * kQueueWords);
    float *fifo = tile + tile_elems;
    float *queue = fifo + kFifo * kFifoWords;
    // per-thread parking slots [word][thread] for path constants that are only needed at shading time (camera weight,
    // splat position, the PCG32 increment): 5 registers less across the traversal loop
    float *park = reinterpret_cast<float *>(smem_raw + A.lens_bytes) + kWarps * (tile_elems + kFifo * kFifoWords + kQueue * kQueueWords) + tid;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.lens);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = tid; i < A.lens_bytes / 16; i += kCtaThreads) dst[i] = src[i];   // header + the n interfaces in use
    }
    const FilmParams &F = A.film;
    const bool box = F.kind == FILTER_BOX;
    const float scale_x = 1.f / (float) F.size_x, scale_y = 1.f / (float) F.size_y;
    const float offs_x = -(float) F.off_x * scale_x, offs_y = -(float) F.off_y * scale_y;
    // per-lane statistics: 32 bits are enough (W*H*spp < 2^32 is spread over the lanes) and leave 6 registers to the traversal
    unsigned n_samples = 0, n_miss = 0, n_rays = 0, n_bounces = 0, n_nodes = 0, n_tris = 0, n_clamp = 0;
    uint2 st[kTravStack];
    // adjoint state: gradient slots shared by the CTA, per-path records (BSDF id, radiance gathered before the bounce)
    __shared__ float s_grad[kAdj ? kGradSlots : 1];
    const bool grad_in_smem = kAdj && A.n_bsdfs * 6 + A.sc.n_emitters * 3 <= kGradSlots;
    if (kAdj) for (int i = tid; i < kGradSlots; i += kCtaThreads) s_grad[i] = 0.f;
    float rec[kAdj ? kMaxRec : 1][4]; int nrec = 0; float3 dL = make_float3(0, 0, 0);
    auto add_grad = [&](int slot, float *gptr, float v) {   // slot: index into s_grad, gptr: the global entry
        if (v == 0.f) return;
        if (grad_in_smem) atomicAdd(&s_grad[slot], v); else atomicAdd(gptr, v);
    };
    __syncthreads();   // lens table staged; from here on the warps of a CTA never wait for each other

    while (true) {
        long long item = 0;
        if (lane == 0) item = (long long) atomicAdd(A.counter, 1ULL);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= A.n_items) break;
        for (int i = lane; i < tile_elems; i += 32) tile[i] = 0.f;
        __syncwarp();
        long long base = item; int sub = -1;
        if (item >= A.n_whole) { const long long k = item - A.n_whole; base = A.n_whole + k / A.n_sub; sub = (int) (k % A.n_sub); }
        const int chunk = (int) (base % A.n_chunks);
        const long long tile_id = (long long) A.shard_index + (base / A.n_chunks) * A.shard_count;
        const int tx = (int) (tile_id % A.tiles_x), ty = (int) (tile_id / A.tiles_x);
        const int px = tx * kTileW + lane, py = ty;                      // the pixel this lane GENERATES samples for
        const bool in_film = px < F.size_x && py < F.size_y;
        const int ox = tx * kTileW + F.border - A.pad, oy = ty * kTileH + F.border - A.pad;   // tensor coords of smem (0,0)
        int j0 = chunk * A.chunk_spp, j1 = min(A.spp, j0 + A.chunk_spp);
        if (sub >= 0) { const int len = max(0, j1 - j0); j1 = j0 + (len * (sub + 1)) / A.n_sub; j0 += (len * sub) / A.n_sub; }
        const int n_rounds = py < F.size_y ? max(0, j1 - j0) : 0;   // generation rounds of this warp
        const uint32_t idx_base = (uint32_t) (py * F.size_x + px) * (uint32_t) A.spp;
        const float posx = (float) (px + F.off_x), posy = (float) (py + F.off_y);
        int round = 0, fifo_head = 0, fifo_count = 0, q_count = 0;       // warp-uniform

        // ---- per-lane path state. phase: 0 = idle, 1 = path segment (closest hit), 2 = shadow ray
        int phase = 0;
        uint64_t rng_state = 0;   // the stream's increment is parked, see `park`
        float3 ro = make_float3(0, 0, 0), rd = make_float3(0, 0, 1);                          // ray being traced now
        float3 wi_w = make_float3(0, 0, 1);                                                   // direction of the path segment
        float3 throughput = make_float3(1, 1, 1), result = make_float3(0, 0, 0), prev_p = make_float3(0, 0, 0);
        float eta = 1.f, prev_pdf = 1.f;
        int depth = 0; bool prev_delta = true, valid_ray = false;
        RayHit seg; seg.t = 0.f; seg.u = seg.v = 0.f; seg.tri = -1;                           // hit of the current segment
        Trav T; trav_init(T, 0.f); T.cur = kTravDone;
        // analytic spheres (kAttr variant only): tested when a ray starts; the BVH traversal then runs with the sphere's
        // distance as its limit, and a shadow ray that a sphere blocks skips the traversal altogether
        // (the sphere's index + 1 rides in T.hit.u, which only an accepted triangle overwrites; its distance is T.tmax; the
        // segment's origin — needed to rebuild a sphere hit in the second half of the bounce — is parked in shared memory)
        auto start_ray = [&](float maxt, bool any_hit) {
            int k = -1;
            if (kAnalytic && A.sc.n_spheres > 0) {
                float t = maxt;
                k = spheres_intersect(A.sc, ro, rd, maxt, any_hit, t);
                if (k >= 0) maxt = t;
            }
            trav_init(T, maxt);
            if (kAnalytic && k >= 0) { T.hit.u = __int_as_float(k + 1); if (any_hit) T.cur = kTravDone; }
        };

        // drain `n` entries from the top of the splat queue (n <= 32), one per lane
        auto drain = [&](int n) {
            __syncwarp();
            if (lane < n) {
                const int s = q_count - n + lane;
                splat_to_tile(F, tile, A.smem_w, ox, oy, queue[0 * kQueue + s], queue[1 * kQueue + s],
                              make_float3(queue[2 * kQueue + s], queue[3 * kQueue + s], queue[4 * kQueue + s]));
            }
            q_count -= n;
            __syncwarp();
        };
        // push one finished sample per lane with `pred`
        auto push_splat = [&](bool pred, float qx, float qy, float3 v) {
            const unsigned m = __ballot_sync(0xffffffffu, pred);
            if (pred) {
                const int s = q_count + __popc(m & lt_mask);
                queue[0 * kQueue + s] = qx; queue[1 * kQueue + s] = qy;
                queue[2 * kQueue + s] = v.x; queue[3 * kQueue + s] = v.y; queue[4 * kQueue + s] = v.z;
            }
            q_count += __popc(m);
        };

        bool exhausted = false;   // no lane active, no camera ray left: only the queue remains to be drained
        while (true) {
            // the only drain site (splat_to_tile is ~500 instructions; inlining it four times cost instruction-cache misses)
            if (q_count >= 32 || (exhausted && q_count > 0)) {
                drain(min(q_count, 32));
                if (q_count >= 32) continue;   // back to this (single) drain site: < 32 entries are left below, 32 + 32 more fit
            }
            if (exhausted) { if (q_count == 0) break; continue; }
            // ---- A. camera rays, 32 at a time: render_sample() draws in the reference's order (integrator.cpp:435-448)
            if (fifo_count < 32 && round < n_rounds) {
                const int j = j0 + round; round++;
                bool hit = false; CameraRay cr; Pcg32 g; float gx = 0.f, gy = 0.f;
                if (in_film) {
                    g.seed_sample(A.seed, idx_base + (uint32_t) j); n_samples++;
                    const float u0 = g.next_f32(), u1 = g.next_f32();
                    const float spx = strict::add(posx, u0), spy = strict::add(posy, u1);
                    const float ax = strict::fmadd(spx, scale_x, offs_x), ay = strict::fmadd(spy, scale_y, offs_y);
                    float a = .5f, b = .5f;
                    if (L.kind == 0) { a = g.next_f32(); b = g.next_f32(); }
                    if (L.shutter_open_time > 0.f) (void) g.next_f32();   // the time sample (integrator.cpp:442-444)
                    cr = sensor_sample_ray(L, L.ifs, ax, ay, a, b);
                    gx = box ? posx : spx; gy = box ? posy : spy;
                    hit = cr.weight > 0.f && A.max_depth != 0;
                    if (!(cr.weight > 0.f)) n_miss++;
                }
                // vignetted (or max_depth == 0): RGB 0, weight 1 (integrator.cpp:461-490)
                if (!kAdj) push_splat(in_film && !hit, gx, gy, make_float3(0, 0, 0));
                const unsigned hm = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int s = (fifo_head + fifo_count + __popc(hm & lt_mask)) & (kFifo - 1);
                    fifo[0 * kFifo + s] = cr.o.x; fifo[1 * kFifo + s] = cr.o.y; fifo[2 * kFifo + s] = cr.o.z;
                    fifo[3 * kFifo + s] = cr.d.x; fifo[4 * kFifo + s] = cr.d.y; fifo[5 * kFifo + s] = cr.d.z;
                    fifo[6 * kFifo + s] = cr.maxt; fifo[7 * kFifo + s] = cr.weight; fifo[8 * kFifo + s] = gx; fifo[9 * kFifo + s] = gy;
                    fifo[10 * kFifo + s] = __uint_as_float((uint32_t) g.state); fifo[11 * kFifo + s] = __uint_as_float((uint32_t) (g.state >> 32));
                    fifo[12 * kFifo + s] = __uint_as_float((uint32_t) g.inc); fifo[13 * kFifo + s] = __uint_as_float((uint32_t) (g.inc >> 32));
                }
                fifo_count += __popc(hm);
                __syncwarp();
            }
            // ---- B. idle lanes take the next camera rays
            {
                const unsigned nm = __ballot_sync(0xffffffffu, phase == 0);
                const int take = min(__popc(nm), fifo_count), rank = __popc(nm & lt_mask);
                if (phase == 0 && rank < take) {
                    const int s = (fifo_head + rank) & (kFifo - 1);
                    ro = make_float3(fifo[0 * kFifo + s], fifo[1 * kFifo + s], fifo[2 * kFifo + s]);
                    rd = make_float3(fifo[3 * kFifo + s], fifo[4 * kFifo + s], fifo[5 * kFifo + s]);
                    start_ray(fifo[6 * kFifo + s], false);
                    const float cam_weight = fifo[7 * kFifo + s], sx = fifo[8 * kFifo + s], sy = fifo[9 * kFifo + s];
                    park[0 * kCtaThreads] = cam_weight; park[1 * kCtaThreads] = sx; park[2 * kCtaThreads] = sy;
                    park[3 * kCtaThreads] = fifo[12 * kFifo + s]; park[4 * kCtaThreads] = fifo[13 * kFifo + s];
                    rng_state = (uint64_t) __float_as_uint(fifo[10 * kFifo + s]) | ((uint64_t) __float_as_uint(fifo[11 * kFifo + s]) << 32);
                    phase = 1;
                    throughput = make_float3(1, 1, 1); result = make_float3(0, 0, 0); eta = 1.f; depth = 0;
                    prev_p = make_float3(0, 0, 0); prev_pdf = 1.f; prev_delta = true;
                    valid_ray = kAnalytic && A.sc.env >= 0 && !A.hide_emitters;     // path.cpp:113-114
                    if (kAdj) {   // transpose of the splat: dL = sum over the footprint of w * d(loss)/d(storage RGB)
                        nrec = 0; dL = make_float3(0, 0, 0);
                        Footprint fp;
                        if (film_footprint(F, sx, sy, fp)) {
                            const int sw = F.size_x + 2 * F.border, shh = F.size_y + 2 * F.border;
                            for (int iy = 0; iy < fp.ny; ++iy) {
                                const int y = fp.y0 + iy;
                                if ((unsigned) y >= (unsigned) shh) continue;
                                for (int ix = 0; ix < fp.nx; ++ix) {
                                    const int x = fp.x0 + ix;
                                    if ((unsigned) x >= (unsigned) sw) continue;
                                    const float4 G = __ldg(reinterpret_cast<const float4 *>(A.grad_storage) + (size_t) y * sw + x);
                                    const float w = fp.wy[iy] * fp.wx[ix];
                                    dL += make_float3(G.x, G.y, G.z) * w;
                                }
                            }
                        }
                        dL *= cam_weight;
                    }
                }
                fifo_head = (fifo_head + take) & (kFifo - 1); fifo_count -= take;
                __syncwarp();
            }
            if (__ballot_sync(0xffffffffu, phase != 0) == 0) {
                exhausted = fifo_count == 0 && round >= n_rounds;
                continue;
            }
            // ---- C. the ONE traversal site: path segments and shadow rays of different lanes advance together
            unsigned work = 0;
            bvh_burst(A.sc.bvh, ro, rd, T, st, phase != 0, phase == 2, A.exit_drop, A.leaf_min, work);
            n_nodes += work & 0xffffu; n_tris += (work >> 16) & 0x7fffu; n_clamp += work >> 31;

            // ---- D. lanes whose ray is finished shade it
            bool done_path = false; float3 value = make_float3(0, 0, 0);
            if (phase != 0 && T.cur == kTravDone) {
                RayHit h = T.hit;
                if (kAnalytic && h.tri < 0 && h.u != 0.f) { h.tri = A.sc.n_tris + __float_as_int(h.u) - 1; h.t = T.tmax; h.u = 0.f; }
                n_rays++;
                bool finish = false, occluded = true;
                Pcg32 rng; rng.state = rng_state;
                rng.inc = (uint64_t) __float_as_uint(park[3 * kCtaThreads]) | ((uint64_t) __float_as_uint(park[4 * kCtaThreads]) << 32);
                const float cam_weight = park[0 * kCtaThreads];
                if (phase == 1) {
                    // ---- first half of one iteration of PathIntegrator::sample's loop (path.cpp:142-208)
                    n_bounces++;
                    seg = h; wi_w = rd;
                    if (kAnalytic) { park[5 * kCtaThreads] = ro.x; park[6 * kCtaThreads] = ro.y; park[7 * kCtaThreads] = ro.z; }
                    const bool hit = h.tri >= 0;
                    bool want_shadow = false;
                    if (kAnalytic && !hit && A.sc.env >= 0 && prev_pdf > 0.f) {
                        // the ray escaped: environment emitter, MIS against the previous bounce's BSDF sample (path.cpp:158-174)
                        const float4 e0 = __ldg(A.sc.emitters + 5 * (size_t) A.sc.env);
                        const float mw = mis_weight(prev_pdf, prev_delta ? 0.f : pdf_environment(A.sc));
                        result += throughput * make_float3(e0.x, e0.y, e0.z) * mw;
                        if (kAdj) {
                            const float3 g = dL * throughput * mw;
                            float *gp = A.grad_emitters + 16 * (size_t) A.sc.env; const int sl = A.n_bsdfs * 6 + A.sc.env * 3;
                            add_grad(sl, gp, g.x); add_grad(sl + 1, gp + 1, g.y); add_grad(sl + 2, gp + 2, g.z);
                        }
                    }
                    if (hit) {
                        valid_ray = true;   // path.cpp:256 uses the loop mask `active`, not `active_next`
                        const Surface si = make_surface<kAttr, kAnalytic>(A.sc, h, ro, rd);
                        const int em = __ldg(A.sc.shape_emitter + si.shape);
                        if (em >= 0) {     // direct emission, MIS against the previous bounce's BSDF sample (:158-174)
                            float3 rel = si.p - prev_p; float dist = sqrtf(dot(rel, rel)); float3 dd = rel * (1.f / dist);
                            float em_pdf = prev_delta ? 0.f : pdf_emitter_direction(A.sc, em, dd, dist, si.sn);
                            if (si.wi.z > 0.f && prev_pdf > 0.f) {
                                float4 e0 = __ldg(A.sc.emitters + 5 * (size_t) em);
                                const float mw = mis_weight(prev_pdf, em_pdf);
                                result += throughput * make_float3(e0.x, e0.y, e0.z) * mw;
                                if (kAdj) {   // radiance enters linearly
                                    const float3 g = dL * throughput * mw;
                                    float *gp = A.grad_emitters + 16 * (size_t) em; const int sl = A.n_bsdfs * 6 + em * 3;
                                    add_grad(sl, gp, g.x); add_grad(sl + 1, gp + 1, g.y); add_grad(sl + 2, gp + 2, g.z);
                                }
                            }
                        }
                        if (depth + 1 < A.max_depth) {
                            finish = true;
                            if (bsdf_is_smooth(A.sc, __ldg(A.sc.shape_bsdf + si.shape))) {
                                // emitter sampling (:193-208) with a COPY of the generator: the six draws of this
                                // iteration are committed when the bounce is finished
                                Pcg32 peek = rng;
                                float e_sx = peek.next_f32(), e_sy = peek.next_f32();
                                EmitterSample es = sample_emitter_direction<kAttr, kAnalytic>(A.sc, si.p, e_sx, e_sy);
                                if (es.pdf != 0.f) {   // shadow ray: Interaction::spawn_ray_to (interaction.h:139-146)
                                    float3 so = offset_p(si.p, si.n, es.p - si.p);
                                    float3 sd = es.p - so; float sdist = sqrtf(dot(sd, sd));
                                    ro = so; rd = sd * (1.f / sdist); start_ray(sdist * (1.f - kShadowEpsilon), true);
                                    want_shadow = true;
                                }
                            }
                        }
                    }
                    if (want_shadow) { phase = 2; finish = false; }
                    else if (!finish) { phase = 0; done_path = true; if (valid_ray) value = result * cam_weight; }
                } else {
                    finish = true; occluded = h.tri >= 0;
                }
                if (finish) {
                    // ---- second half (path.cpp:210-281): BSDF eval/sample, NEE contribution, next ray, Russian roulette
                    const float3 seg_o = kAnalytic ? make_float3(park[5 * kCtaThreads], park[6 * kCtaThreads], park[7 * kCtaThreads]) : make_float3(0, 0, 0);
                    const Surface si = make_surface<kAttr, kAnalytic>(A.sc, seg, seg_o, wi_w);
                    const int bsdf_id = __ldg(A.sc.shape_bsdf + si.shape);
                    float e_sx = rng.next_f32(), e_sy = rng.next_f32();
                    float s1 = rng.next_f32(), s2x = rng.next_f32(), s2y = rng.next_f32();
                    float rr_sample = rng.next_f32();
                    EmitterSample es; es.pdf = 0.f; float3 wo = make_float3(0, 0, 0);
                    const bool active_em = phase == 2 && !occluded;
                    if (active_em) { es = sample_emitter_direction<kAttr, kAnalytic>(A.sc, si.p, e_sx, e_sy); wo = si.to_local(es.d); }
                    BsdfResult bs = bsdf_eval_pdf_sample<kAttr>(A.sc, bsdf_id, si.uv, si.wi, wo, s1, s2x, s2y);
                    if (kAdj) {   // everything gathered from here on is linear in this surface's (specular) reflectance
                        // the sampled lobe's colour: (specular) reflectance, or the dielectric's specular_transmittance
                        if (nrec < kMaxRec && __ldg(reinterpret_cast<const float *>(A.sc.bsdfs) + 16 * (size_t) bsdf_id + 15) == 0.f) {   // (textured reflectances are not parameters)
                            rec[nrec][0] = __int_as_float(2 * bsdf_id + (bs.transmitted ? 1 : 0)); rec[nrec][1] = result.x; rec[nrec][2] = result.y; rec[nrec][3] = result.z; nrec++;
                        }
                    }
                    if (active_em) {
                        const float3 c3 = throughput * (bs.val * es.weight * mis_weight(es.pdf, bs.pdf));
                        result += c3;
                        if (kAdj) {
                            const int cnt = A.sc.n_emitters, ei = cnt > 1 ? min((int) (e_sx * (float) cnt), cnt - 1) : 0;
                            const float4 e0 = __ldg(A.sc.emitters + 5 * (size_t) ei);
                            float *gp = A.grad_emitters + 16 * (size_t) ei; const int sl = A.n_bsdfs * 6 + ei * 3;
                            if (e0.x != 0.f) add_grad(sl, gp, dL.x * c3.x / e0.x);
                            if (e0.y != 0.f) add_grad(sl + 1, gp + 1, dL.y * c3.y / e0.y);
                            if (e0.z != 0.f) add_grad(sl + 2, gp + 2, dL.z * c3.z / e0.z);
                        }
                    }
                    float3 wo_w = si.to_world(bs.wo);
                    ro = offset_p(si.p, si.n, wo_w); rd = wo_w; start_ray(kLargest, false);
                    throughput *= bs.weight; eta *= bs.seta;
                    valid_ray = true;
                    prev_p = si.p; prev_pdf = bs.spdf; prev_delta = bs.delta;
                    depth++;
                    float tmax = max3(throughput);
                    float rr_prob = fminf(tmax * eta * eta, .95f);
                    bool rr_active = depth >= A.rr_depth;
                    if (rr_active) throughput *= 1.f / rr_prob;
                    bool alive = (!rr_active || rr_sample < rr_prob) && tmax != 0.f;
                    phase = alive ? 1 : 0;
                    if (!alive) { done_path = true; value = result * cam_weight; T.cur = kTravDone; }
                    rng_state = rng.state;
                }
            }
            if (!kAdj) push_splat(done_path, park[1 * kCtaThreads], park[2 * kCtaThreads], value);
            else if (done_path) {   // suffix of the path behind every recorded bounce, divided by that bounce's reflectance
                for (int k = 0; k < nrec; ++k) {
                    const int code = __float_as_int(rec[k][0]), b = code >> 1, col = (code & 1) ? 11 : 2;
                    const float *row = reinterpret_cast<const float *>(A.sc.bsdfs) + 16 * (size_t) b + col;
                    float *gp = A.grad_bsdfs + 16 * (size_t) b + col; const int sl = b * 6 + ((code & 1) ? 3 : 0);
                    const float r0 = __ldg(row), r1 = __ldg(row + 1), r2 = __ldg(row + 2);
                    if (r0 != 0.f) add_grad(sl, gp, dL.x * (result.x - rec[k][1]) / r0);
                    if (r1 != 0.f) add_grad(sl + 1, gp + 1, dL.y * (result.y - rec[k][2]) / r1);
                    if (r2 != 0.f) add_grad(sl + 2, gp + 2, dL.z * (result.z - rec[k][3]) / r2);
                }
                nrec = 0;
            }
        }
        // ---- drain what is left, then put_block(): flush the tile into the film accumulator
        if (kAdj) { __syncwarp(); continue; }
        const int sw = F.size_x + 2 * F.border, shh = F.size_y + 2 * F.border;
        for (int i = lane; i < A.smem_w * A.smem_h; i += 32) {
            int ly = i / A.smem_w, lx = i - ly * A.smem_w;
            int x = ox + lx, y = oy + ly;
            if ((unsigned) x >= (unsigned) sw || (unsigned) y >= (unsigned) shh) continue;
            float4 v = reinterpret_cast<const float4 *>(tile)[i];
            if (v.w == 0.f && v.x == 0.f && v.y == 0.f && v.z == 0.f) continue;
            // one 128-bit reduction per pixel (sm_90+: atomicAdd on float4, element-wise atomic) instead of four scalar ones
            atomicAdd(reinterpret_cast<float4 *>(A.accum) + ((size_t) y * sw + x), v);
        }
        __syncwarp();
    }
    if (kAdj && grad_in_smem) {
        __syncthreads();
        for (int i = tid; i < A.n_bsdfs * 6 + A.sc.n_emitters * 3; i += kCtaThreads) {
            const float v = s_grad[i];
            if (v == 0.f) continue;
            float *dst;
            if (i < A.n_bsdfs * 6) { const int bb = i / 6, c = i - 6 * bb; dst = A.grad_bsdfs + 16 * (size_t) bb + (c < 3 ? 2 + c : 8 + c); }
            else { const int j = i - A.n_bsdfs * 6, e = j / 3; dst = A.grad_emitters + 16 * (size_t) e + (j - 3 * e); }
            atomicAdd(dst, v);
        }
    }
    if (A.stats) {
        unsigned long long c[7] = { n_samples, n_miss, n_rays, n_bounces, n_nodes, n_tris, n_clamp };
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xffffffffu, c[k], o);
            if (lane == 0) atomicAdd(A.stats + k, c[k]);
        }
    }
}

// unit-level Scene::ray_intersect / ray_test
__global__ void __launch_bounds__(128) k_intersect(SceneView sc, int any_hit, int64_t n, const float *__restrict__ od7, float *__restrict__ out4) {
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float *r = od7 + 7 * i;
        float3 o = make_float3(r[0], r[1], r[2]), d = make_float3(r[3], r[4], r[5]);
        unsigned work = 0;
        RayHit h = scene_intersect(sc, o, d, r[6], any_hit != 0, work);
        float *q = out4 + 4 * i;
        q[0] = h.t; q[1] = h.u; q[2] = h.v;
        if (h.tri >= sc.n_tris) q[3] = (float) (-2 - (h.tri - sc.n_tris));                 // analytic sphere k: -2 - k
        else q[3] = h.tri >= 0 ? (float) __float_as_int(__ldg(sc.bvh.tris + 3 * (size_t) h.tri + 2).z) : -1.f;
    }
}

}  // namespace lrt

// ================================================================================ host side
namespace lrt {

int launch_render(const lrt_scene *sc, lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                  float *accum_dev, const float *grad_storage, float *grad_bsdfs, float *grad_emitters, cudaStream_t stream) {
    RenderArgs A; memset(&A, 0, sizeof A);
    int rc = film_params_from_desc(film, A.film); if (rc) return rc;
    LRT_REQUIRE(film->channels == 4, "render needs the R,G,B,W film of hdrfilm (channels = 4)");
    LRT_REQUIRE(rp->spp > 0 && rp->max_depth >= -1 && rp->shard_count >= 1 && rp->shard_index >= 0 && rp->shard_index < rp->shard_count, "bad render parameters");
    // one PCG32 stream per sample: the wavefront must fit 32 bits (integrator.cpp:247-265 splits into passes)
    LRT_REQUIRE((double) film->size_x * film->size_y * rp->spp < 4294967296.0, "W*H*spp exceeds 2^32: render in passes");
    rc = sensor->upload(stream); if (rc) return rc;
    A.sc = sc->view; A.lens = sensor->dev;
    A.seed = rp->seed; A.spp = rp->spp; A.max_depth = rp->max_depth < 0 ? 0x7fffffff : rp->max_depth; A.rr_depth = rp->rr_depth;
    A.hide_emitters = rp->hide_emitters;
    A.film.coalesce = rp->spp >= 4 ? film->coalesce : 0;   // integrator.cpp:290
    const FilmParams &F = A.film;
    A.pad = F.kind == FILTER_BOX ? 0 : (F.coalesce ? F.n_half : (int) std::ceil(F.radius + .5f));
    // +1: pos + u (u < 1) can round up to the next integer in float32 (ulp(2048) = 2.4e-4), shifting the window by one
    A.smem_w = kTileW + 2 * A.pad + 1; A.smem_h = kTileH + 2 * A.pad + 1;
    A.tiles_x = (F.size_x + kTileW - 1) / kTileW;
    int tiles_y = (F.size_y + kTileH - 1) / kTileH;
    long long n_tiles = (long long) A.tiles_x * tiles_y;
    long long my_tiles = (n_tiles - rp->shard_index + rp->shard_count - 1) / rp->shard_count;
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // only the header and the interfaces in use are staged (not all 32 records, not the exit-pupil LUT): every KB of
    // shared memory given back is L1 for the traversal
    A.lens_bytes = (int) ((offsetof(LensSystem, ifs) + (size_t) sensor->host.n * sizeof(LensIf) + 15) / 16 * 16);
    const bool attr = sc->view.tri_attr != nullptr || sc->view.textures != nullptr || sc->view.em_tris != nullptr || sc->view.n_spheres > 0 ||
                      sc->view.env >= 0;
    const bool analytic = sc->view.n_spheres > 0 || sc->view.env >= 0;
    size_t smem = (size_t) A.lens_bytes + (size_t) kCtaThreads * (analytic ? 8 : 5) * 4 + (size_t) kWarps * ((size_t) A.smem_w * A.smem_h * 4 + kFifo * kFifoWords + kQueue * kQueueWords) * 4;
    {   // traversal bursts end when this many 32nds of their lanes have finished (LRT_EXIT_DROP overrides; tuning knob)
        const char *e = getenv("LRT_EXIT_DROP"); A.exit_drop = e ? atoi(e) : 26;
        if (A.exit_drop < 1) A.exit_drop = 1; if (A.exit_drop > 32) A.exit_drop = 32;
        e = getenv("LRT_LEAF_MIN"); A.leaf_min = e ? atoi(e) : 6;
        if (A.leaf_min < 1) A.leaf_min = 1;
    }
    // wide reconstruction filters make the private film tiles large; beyond the opt-in limit the render is refused
    LRT_REQUIRE(smem <= 200 * 1024, "reconstruction filter radius too large for the shared-memory film tiles of this build");
    {   // the shared-memory opt-in is a per-device function attribute
        static std::mutex mu; static bool attr_set[64] = {};
        std::lock_guard<std::mutex> lock(mu);
        if (dev >= 64 || !attr_set[dev]) {
            void (*all[])(RenderArgs) = { k_render<false, 0>, k_render<false, 1>, k_render<false, 2>, k_render<true, 0>, k_render<true, 1>, k_render<true, 2> };
            for (auto f : all) LRT_CUDA(cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            if (dev < 64) attr_set[dev] = true;
        }
    }
    const bool adjoint = grad_storage != nullptr;
    // the one-pass adjoint keeps one record per bounce of a path: longer paths are refused, not truncated
    LRT_REQUIRE(!adjoint || (rp->max_depth >= 0 && rp->max_depth <= kMaxRec + 1),
                "lrt_render_backward: max_depth must be in [0, 17] (one gradient record per bounce; unbounded paths are not supported)");
    A.grad_storage = grad_storage; A.grad_bsdfs = grad_bsdfs; A.grad_emitters = grad_emitters; A.n_bsdfs = sc->n_bsdfs;
    int per_sm = 1;
    void (*kern)(RenderArgs) = adjoint ? (analytic ? k_render<true, 2> : attr ? k_render<true, 1> : k_render<true, 0>)
                                       : (analytic ? k_render<false, 2> : attr ? k_render<false, 1> : k_render<false, 0>);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kCtaThreads, smem);
    if (per_sm < 1) per_sm = 1;
    long long resident = (long long) sms * per_sm;
    // split the samples of a pixel into chunks until there are ~4 work items per resident warp
    int n_chunks = 1;
    while (my_tiles * n_chunks < 4 * resident * kWarps && n_chunks * 2 <= rp->spp) n_chunks *= 2;
    if (rp->chunks > 0) n_chunks = std::min(rp->chunks, rp->spp);
    A.n_chunks = n_chunks; A.chunk_spp = (rp->spp + n_chunks - 1) / n_chunks;
    A.shard_index = rp->shard_index; A.shard_count = rp->shard_count;
    // guided tail: the last (resident warps) base items are handed out in pieces of >= 4 spp (LRT_TAIL_SPLIT=0 turns it off;
    // an explicit rp->chunks does too). Those pieces are ~1/16 of a full item: one resident warp's worth of them keeps every
    // warp busy while the last full-size items drain.
    {
        const long long n_base = my_tiles * n_chunks;
        const char *e = getenv("LRT_TAIL_SPLIT");
        int n_sub = std::min(16, A.chunk_spp / 4);
        long long n_split = std::min(n_base, resident * kWarps);
        if (rp->chunks > 0 || n_sub < 2 || (e && atoi(e) == 0)) { n_sub = 1; n_split = 0; }
        A.n_sub = n_sub; A.n_whole = n_base - n_split; A.n_items = A.n_whole + n_split * n_sub;
    }
    // every launch takes its own work counter from a ring (renders of one scene may overlap on different streams); the
    // statistics block is per scene: it describes the last render launched and is meaningless while two overlap
    unsigned long long *ctr = sc->d_counter + (sc->launch_seq.fetch_add(1, std::memory_order_relaxed) % kCounterRing);
    A.counter = ctr; A.stats = sc->d_counter + kCounterRing; A.accum = accum_dev;
    LRT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(unsigned long long), stream));
    LRT_CUDA(cudaMemsetAsync(A.stats, 0, 8 * sizeof(unsigned long long), stream));
    long long grid = std::min(resident, (A.n_items + kWarps - 1) / kWarps);
    if (grid < 1) return LRT_OK;
    kern<<<(int) grid, kCtaThreads, smem, stream>>>(A);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

}  // namespace lrt

extern "C" int lrt_scene_intersect(const lrt_scene *sc, int32_t any_hit, int64_t n, const float *od7_dev, float *out4_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && n >= 0 && (n == 0 || (od7_dev && out4_dev)), "bad argument");   // an empty batch may come with null pointers
    if (n == 0) return LRT_OK;
    int64_t g = (n + 127) / 128; if (g > 148 * 32) g = 148 * 32;
    lrt::k_intersect<<<(int) g, 128, 0, (cudaStream_t) stream>>>(sc->view, any_hit, n, od7_dev, out4_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}
