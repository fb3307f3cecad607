// lensrt — the reference's SCALAR render loop as a kernel (lrt_render_params.rng_mode = 1): a parity mode, not the fast path.
// Replaces SamplingIntegrator::render, scalar branch (src/render/integrator.cpp:163-245) + render_block (:369-418) +
// render_sample (:419-491) with PathIntegrator::sample in its scalar control flow (src/integrators/path.cpp:95-287:
// `break` before any draw once the path cannot continue, emitter samples drawn only at smooth surfaces) and the scalar
// ImageBlock::put (src/render/imageblock.cpp:264-379: non-coalesced window, 31-tap discretised filter).
//
// Sampler convention (integrator.cpp:195-200,388-402; sampler.cpp:131): ONE PCG32 stream per pixel,
//   rng.seed(1, base_seed + seed * W * H + block_id * block_size^2 + morton(x, y), PCG32_DEFAULT_STREAM),
// block_id = position of the pixel's block along the spiral (src/render/spiral.cpp:26-75), the spp samples of the pixel
// consecutive draws of that stream. A pixel's samples are therefore sequential: one thread owns one pixel. Images of
// this mode can be compared sample for sample with the reference's scalar_rgb variant (tests/golden/ref_renders.npz);
// the default mode (rng_mode 0, lrt_render.cu) follows the JIT variants' per-sample streams.
#include "lrt_internal.h"
#include "lrt_rng.cuh"
#include "lrt_shade.cuh"
#include "lrt_scene.h"
#include <vector>

namespace lrt {

struct ScalarArgs {
    SceneView sc;
    const LensSystem *lens;
    FilmParams film;
    uint32_t seed0;                  // base_seed + seed * W * H
    int spp, max_depth, rr_depth, block_size, blocks_x, hide_emitters;
    const uint32_t *block_id;        // [blocks_y * blocks_x] spiral position of every block
    float *accum;
    unsigned long long *stats;       // [0] samples, [1] lens misses, [2] rays, [3] bounces
};

LRT_D uint32_t spread_bits(uint32_t x) {   // 16 low bits -> even bit positions (dr::morton_encode)
    x &= 0xffffu; x = (x | (x << 8)) & 0x00ff00ffu; x = (x | (x << 4)) & 0x0f0f0f0fu;
    x = (x | (x << 2)) & 0x33333333u; x = (x | (x << 1)) & 0x55555555u;
    return x;
}

// PathIntegrator::sample, scalar control flow. Dead lens rays (weight 0) are traced as well: the reference does not mask
// them and they consume random numbers of the pixel's stream.
LRT_D float3 path_sample_scalar(const ScalarArgs &A, Pcg32 &rng, float3 ro, float3 rd, float maxt, unsigned &n_rays, unsigned &n_bounces) {
    float3 throughput = make_float3(1, 1, 1), result = make_float3(0, 0, 0), prev_p = make_float3(0, 0, 0);
    float eta = 1.f, prev_pdf = 1.f; bool prev_delta = true, valid_ray = false;
    int depth = 0;
    if (A.max_depth == 0) return result;
    valid_ray = A.sc.env >= 0 && !A.hide_emitters;                   // path.cpp:113-114
    while (true) {
        unsigned work = 0;
        const RayHit h = scene_intersect(A.sc, ro, rd, maxt, false, work);
        n_rays++; n_bounces++;
        if (h.tri < 0) {     // the ray escaped: the environment emitter, MIS against the previous bounce (path.cpp:158-174)
            if (A.sc.env >= 0 && prev_pdf > 0.f) {
                const float4 e0 = __ldg(A.sc.emitters + 5 * (size_t) A.sc.env);
                result += throughput * make_float3(e0.x, e0.y, e0.z) * mis_weight(prev_pdf, prev_delta ? 0.f : pdf_environment(A.sc));
            }
            break;
        }
        const Surface si = make_surface(A.sc, h, ro, rd);
        const int em = __ldg(A.sc.shape_emitter + si.shape);
        if (em >= 0) {   // direct emission, MIS against the previous bounce's BSDF sample (path.cpp:158-174)
            const float3 rel = si.p - prev_p; const float dist = sqrtf(dot(rel, rel)); const float3 dd = rel * (1.f / dist);
            const float em_pdf = prev_delta ? 0.f : pdf_emitter_direction(A.sc, em, dd, dist, si.sn);
            if (si.wi.z > 0.f && prev_pdf > 0.f) {
                const float4 e0 = __ldg(A.sc.emitters + 5 * (size_t) em);
                result += throughput * make_float3(e0.x, e0.y, e0.z) * mis_weight(prev_pdf, em_pdf);
            }
        }
        if (!(depth + 1 < A.max_depth)) break;                       // path.cpp:176-179
        const int bsdf_id = __ldg(A.sc.shape_bsdf + si.shape);
        EmitterSample es; es.pdf = 0.f; float3 wo = make_float3(0, 0, 0);
        bool active_em = bsdf_is_smooth(A.sc, bsdf_id);
        if (active_em) {                                             // scene.cpp:299-353 with test_visibility = true
            const float e_sx = rng.next_f32(), e_sy = rng.next_f32();
            es = sample_emitter_direction(A.sc, si.p, e_sx, e_sy);
            if (es.pdf != 0.f) {
                const float3 so = offset_p(si.p, si.n, es.p - si.p);
                const float3 sd = es.p - so; const float sdist = sqrtf(dot(sd, sd));
                unsigned w2 = 0;
                const RayHit sh = scene_intersect(A.sc, so, sd * (1.f / sdist), sdist * (1.f - kShadowEpsilon), true, w2);
                n_rays++;
                if (sh.tri >= 0) es.pdf = 0.f;
            }
            active_em = es.pdf != 0.f;
            wo = si.to_local(es.d);
        }
        const float s1 = rng.next_f32(), s2x = rng.next_f32(), s2y = rng.next_f32();
        const BsdfResult bs = bsdf_eval_pdf_sample(A.sc, bsdf_id, si.uv, si.wi, wo, s1, s2x, s2y);
        if (active_em) result += throughput * (bs.val * es.weight * mis_weight(es.pdf, bs.pdf));
        const float3 wo_w = si.to_world(bs.wo);
        ro = offset_p(si.p, si.n, wo_w); rd = wo_w; maxt = kLargest;
        throughput *= bs.weight; eta *= bs.seta;
        valid_ray = true;
        prev_p = si.p; prev_pdf = bs.spdf; prev_delta = bs.delta;
        depth++;
        const float tmax = max3(throughput);
        const float rr_prob = fminf(tmax * eta * eta, .95f);
        const bool rr_active = depth >= A.rr_depth, rr_continue = rng.next_f32() < rr_prob;
        if (rr_active) throughput *= 1.f / rr_prob;
        if (!((!rr_active || rr_continue) && tmax != 0.f)) break;
    }
    return valid_ray ? result : make_float3(0, 0, 0);
}

__global__ void __launch_bounds__(64) k_render_scalar(ScalarArgs A) {
    const FilmParams &F = A.film;
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned n_samples = 0, n_miss = 0, n_rays = 0, n_bounces = 0;
    if (pix < F.size_x * F.size_y) {
        const int py = pix / F.size_x, px = pix - py * F.size_x;
        const int bs = A.block_size, bx = px / bs, by = py / bs;
        const uint32_t morton = spread_bits((uint32_t) (px - bx * bs)) | (spread_bits((uint32_t) (py - by * bs)) << 1);
        Pcg32 rng;
        rng.seed((uint64_t) (A.seed0 + __ldg(A.block_id + by * A.blocks_x + bx) * (uint32_t) (bs * bs) + morton), 0xda3e39cb94b95bdbULL);
        const LensSystem &L = *A.lens;
        const bool box = F.kind == FILTER_BOX;
        const float scale_x = 1.f / (float) F.size_x, scale_y = 1.f / (float) F.size_y;
        const float offs_x = -(float) F.off_x * scale_x, offs_y = -(float) F.off_y * scale_y;
        const float posx = (float) (px + F.off_x), posy = (float) (py + F.off_y);
        const int sw = F.size_x + 2 * F.border, shh = F.size_y + 2 * F.border;
        for (int j = 0; j < A.spp; ++j) {
            n_samples++;
            const float u0 = rng.next_f32(), u1 = rng.next_f32();
            const float spx = strict::add(posx, u0), spy = strict::add(posy, u1);
            const float ax = strict::fmadd(spx, scale_x, offs_x), ay = strict::fmadd(spy, scale_y, offs_y);
            float a = .5f, b = .5f;
            if (L.kind == 0) { a = rng.next_f32(); b = rng.next_f32(); }
            if (L.shutter_open_time > 0.f) (void) rng.next_f32();
            const CameraRay cr = sensor_sample_ray(L, L.ifs, ax, ay, a, b);
            if (!(cr.weight > 0.f)) n_miss++;
            const float3 Lr = path_sample_scalar(A, rng, cr.o, cr.d, cr.maxt, n_rays, n_bounces);
            const float3 value = cr.weight > 0.f ? Lr * cr.weight : make_float3(0, 0, 0);
            Footprint fp;
            if (!film_footprint<true>(F, box ? posx : spx, box ? posy : spy, fp)) continue;
            for (int iy = 0; iy < fp.ny; ++iy) {
                const int y = fp.y0 + iy;
                if ((unsigned) y >= (unsigned) shh) continue;
                for (int ix = 0; ix < fp.nx; ++ix) {
                    const int x = fp.x0 + ix;
                    if ((unsigned) x >= (unsigned) sw) continue;
                    const float w = fp.wx[ix] * fp.wy[iy];
                    float *dst = A.accum + ((size_t) y * sw + x) * 4;
                    atomicAdd(dst, value.x * w); atomicAdd(dst + 1, value.y * w); atomicAdd(dst + 2, value.z * w); atomicAdd(dst + 3, w);
                }
            }
        }
    }
    if (A.stats) {
        unsigned long long c[4] = { n_samples, n_miss, n_rays, n_bounces };
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xffffffffu, c[k], o);
            if ((threadIdx.x & 31) == 0 && c[k]) atomicAdd(A.stats + k, c[k]);
        }
    }
}

// Spiral::next_block order (spiral.cpp:26-75): block (x, y) -> its position along the spiral
static std::vector<uint32_t> spiral_ids(int bx, int by) {
    std::vector<uint32_t> id((size_t) bx * by, 0u);
    int px = bx / 2, py = by / 2, dir = 0, steps_left = 1, spiral_size = 1;   // Right, Down, Left, Up
    const int n = bx * by;
    for (int counter = 0; counter < n; ++counter) {
        id[(size_t) py * bx + px] = (uint32_t) counter;
        if (counter + 1 == n) break;
        do {
            switch (dir) { case 0: ++px; break; case 1: ++py; break; case 2: --px; break; default: --py; break; }
            if (--steps_left == 0) {
                dir = (dir + 1) % 4;
                if (dir == 2 || dir == 0) ++spiral_size;
                steps_left = spiral_size;
            }
        } while (px < 0 || py < 0 || px >= bx || py >= by);
    }
    return id;
}

int launch_render_scalar(const lrt_scene *sc, lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                         float *accum_dev, cudaStream_t stream) {
    ScalarArgs A; memset(&A, 0, sizeof A);
    int rc = film_params_from_desc(film, A.film); if (rc) return rc;
    LRT_REQUIRE(film->channels == 4, "render needs the R,G,B,W film of hdrfilm (channels = 4)");
    LRT_REQUIRE(rp->spp > 0 && rp->max_depth >= -1, "bad render parameters");
    LRT_REQUIRE(rp->shard_count == 1 && rp->shard_index == 0, "the scalar sampler convention (rng_mode 1) is a single-GPU parity mode");
    LRT_REQUIRE((long long) film->size_x * film->size_y < (1ll << 31), "film too large");
    rc = sensor->upload(stream); if (rc) return rc;
    A.film.coalesce = 0; A.film.discretized = 1;       // imageblock.cpp:264-379 with JIT = false
    A.sc = sc->view; A.lens = sensor->dev;
    A.block_size = rp->block_size > 0 ? rp->block_size : 32;     // MI_BLOCK_SIZE (integrator.cpp:168-179)
    LRT_REQUIRE((A.block_size & (A.block_size - 1)) == 0 && A.block_size <= 1024, "block_size must be a power of two");
    A.blocks_x = (film->size_x + A.block_size - 1) / A.block_size;
    const int blocks_y = (film->size_y + A.block_size - 1) / A.block_size;
    A.seed0 = rp->base_seed + rp->seed * (uint32_t) (film->size_x * film->size_y);   // integrator.cpp:196, sampler.cpp:119
    A.spp = rp->spp; A.max_depth = rp->max_depth < 0 ? 0x7fffffff : rp->max_depth; A.rr_depth = rp->rr_depth;
    A.hide_emitters = rp->hide_emitters;
    const std::vector<uint32_t> ids = spiral_ids(A.blocks_x, blocks_y);
    uint32_t *d_ids = nullptr;
    LRT_CUDA(cudaMallocAsync((void **) &d_ids, ids.size() * sizeof(uint32_t), stream));
    LRT_CUDA(cudaMemcpyAsync(d_ids, ids.data(), ids.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    LRT_CUDA(cudaStreamSynchronize(stream));           // `ids` is pageable host memory
    A.block_id = d_ids; A.accum = accum_dev; A.stats = sc->d_counter + kCounterRing;
    LRT_CUDA(cudaMemsetAsync(A.stats, 0, 8 * sizeof(unsigned long long), stream));
    const int npix = film->size_x * film->size_y;
    k_render_scalar<<<(npix + 63) / 64, 64, 0, stream>>>(A);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    LRT_CUDA(cudaFreeAsync(d_ids, stream));
    return LRT_OK;
}

}  // namespace lrt
