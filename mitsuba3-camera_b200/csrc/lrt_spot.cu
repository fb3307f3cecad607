// lensrt — the lens-design hot loop: light-traced spot images through the realistic lens and their adjoint.
//
// Forward = what the reference's `ptracer` does when it connects an emitter vertex to the fork's sensor
// (src/integrators/ptracer.cpp connect_sensor -> RealisticLensCamera::sample_direction,
// src/sensors/realisticlens.cpp:2301-2394 -> ImageBlock::put with normalize=true, coalesce=false,
// src/render/integrator.cpp:706-717, sample_scale = 1/spp :671-676): for every sample draw a point on the
// front element, trace the ray source -> front -> ... -> rear -> film plane, splat
// intensity * weight / spp at ds.uv. A distant point source stands in for the notebooks' directional field
// sources (notebooks/scripts/lens_design/design_problem.py:46-199).
//
// Adjoint = d(loss)/d(interface table) given d(loss)/d(film accumulator): the same samples are replayed
// (same PCG32 streams, same masks), the splat is transposed into an adjoint of the splat POSITION (filter
// weight derivative, the smooth "where does the light land" term the notebooks rely on) and of the sample
// weight, and lrt_lens_adj.cuh sweeps the interfaces in reverse. No tape beyond 6 floats per interface.
//
// Kernel shape: persistent CTAs of 128 threads, lens table in shared memory, sample idx -> lane by a
// grid-stride loop (idx = shard + k * shards, so any GPU count replays the same streams). Spot images put
// nearly all samples into a few pixels, so splats go to a per-CTA shared-memory window centred on the chief
// ray (shared float atomics) and only the window's non-zero pixels reach global memory; samples that fall
// outside use global atomics. The adjoint accumulates parameter gradients per thread in local memory,
// then warp shuffle -> shared -> one global atomic per CTA and slot.
#include "lrt_internal.h"
#include "lrt_rng.cuh"
#include "lrt_lens_adj.cuh"

namespace lrt {

constexpr int kSpotThreads = 128;
constexpr int kWin = 32;            // shared window: kWin x kWin pixels x 4 channels = 16 KB
constexpr int kSlots = 12;          // per-interface gradient slots kept per thread: z, R|c, K, a0..a6, ior_l, ior_r

struct SpotArgs {
    const LensSystem *lens;
    FilmParams film;
    uint32_t seed; int spp;
    float src[3], intensity[3], origin_offset;
    long long n_my; int shard_index, shard_count;
    int win_x0, win_y0;             // tensor coordinates of the window origin
    float scale;
    float *accum;                   // forward: [sy, sx, 4]
    const float *grad_accum;        // adjoint: same shape
    float *grad_rows;               // adjoint: [n_if, 16]
    unsigned long long *stats;      // [0] samples, [1] valid
};

// derivative of the reconstruction filter (same polynomial / tent as filter_eval)
LRT_HD float filter_eval_d(const FilmParams &F, float x) {
    if (F.kind == FILTER_GAUSSIAN) {
        const float *c = F.coeff;
        float x2 = x * x, P = 0.f, Pp = 0.f;
        for (int i = 9; i >= 0; --i) { Pp = Pp * x2 + P; P = P * x2 + c[i]; }
        return P > 0.f ? 2.f * x * Pp : 0.f;
    }
    if (F.kind == FILTER_TENT) {
        float ax = fabsf(x * F.inv_radius);
        return (ax < 1.f && x != 0.f) ? (x > 0.f ? -F.inv_radius : F.inv_radius) : 0.f;
    }
    return 0.f;
}

// Non-coalesced footprint (imageblock.cpp:264-412) with per-axis NORMALISED weights nx, ny (W = nx_i ny_j) and
// their derivatives with respect to the sample position. Returns false if the footprint is empty.
struct FootprintD { int x0, y0, nx, ny; float wx[kMaxTaps], wy[kMaxTaps], dwx[kMaxTaps], dwy[kMaxTaps]; };

LRT_HD bool spot_footprint(const FilmParams &F, float px, float py, FootprintD &fp, bool want_d) {
    int sx = F.size_x + 2 * F.border, sy = F.size_y + 2 * F.border;
    float pfx = px + ((float) (F.border - F.off_x) - .5f), pfy = py + ((float) (F.border - F.off_y) - .5f);
    float p0x = pfx - F.radius, p0y = pfy - F.radius;
    int cx = (int) ceilf(p0x), cy = (int) ceilf(p0y);
    int x0 = max(cx, 0), y0 = max(cy, 0);
    int x1 = min((int) floorf(pfx + F.radius), sx - 1), y1 = min((int) floorf(pfy + F.radius), sy - 1);
    if (x0 > x1 || y0 > y1) return false;
    fp.x0 = x0; fp.y0 = y0; fp.nx = min(x1 - x0 + 1, F.count_max); fp.ny = min(y1 - y0 + 1, F.count_max);
    float a = 1.f, b = 1.f, da = 0.f, db = 0.f;
    if (F.normalize) {   // :318-342: sums start at the unclipped window
        a = b = 0.f;
        float rx = (float) cx - pfx, ry = (float) cy - pfy;
        for (int i = 0; i < F.count_max; ++i) {
            a += filter_eval(F, rx); b += filter_eval(F, ry);
            if (want_d) { da -= filter_eval_d(F, rx); db -= filter_eval_d(F, ry); }
            rx += 1.f; ry += 1.f;
        }
    }
    float ia = a != 0.f ? 1.f / a : 0.f, ib = b != 0.f ? 1.f / b : 0.f;
    if (F.normalize && (a == 0.f || b == 0.f)) { ia = 0.f; ib = 0.f; }
    float rx = (float) x0 - pfx, ry = (float) y0 - pfy;
    for (int i = 0; i < F.count_max; ++i) {
        float fx = filter_eval(F, rx), fy = filter_eval(F, ry);
        fp.wx[i] = fx * ia; fp.wy[i] = fy * ib;
        if (want_d) {
            fp.dwx[i] = (-filter_eval_d(F, rx) - fp.wx[i] * da) * ia;
            fp.dwy[i] = (-filter_eval_d(F, ry) - fp.wy[i] * db) * ib;
        }
        rx += 1.f; ry += 1.f;
    }
    return true;
}

template <bool kAdjoint>
__global__ void __launch_bounds__(kSpotThreads) k_spot(SpotArgs A) {
    __shared__ LensSystem L;
    __shared__ float win[kAdjoint ? 1 : kWin * kWin * 4];
    __shared__ float s_grad[kAdjoint ? kMaxInterfaces * kSlots : 1];
    const int tid = threadIdx.x, lane = tid & 31;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.lens);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = tid; i < (int) (sizeof(LensSystem) / 16); i += kSpotThreads) dst[i] = src[i];
    }
    if (!kAdjoint) for (int i = tid; i < kWin * kWin * 4; i += kSpotThreads) win[i] = 0.f;
    if (kAdjoint) for (int i = tid; i < kMaxInterfaces * kSlots; i += kSpotThreads) s_grad[i] = 0.f;
    __syncthreads();
    const FilmParams &F = A.film;
    const int sw = F.size_x + 2 * F.border, sh = F.size_y + 2 * F.border;
    const strict::V3 src_p = strict::v3(A.src[0], A.src[1], A.src[2]);
    const int n_if = L.n;
    unsigned long long n_samples = 0, n_valid = 0;
    float g_loc[kAdjoint ? kMaxInterfaces : 1][kSlots];
    if (kAdjoint) for (int i = 0; i < n_if; ++i) for (int k = 0; k < kSlots; ++k) g_loc[i][k] = 0.f;

    for (long long i = (long long) blockIdx.x * kSpotThreads + tid; i < A.n_my; i += (long long) gridDim.x * kSpotThreads) {
        const uint32_t idx = (uint32_t) (A.shard_index + i * A.shard_count);
        Pcg32 rng; rng.seed_sample(A.seed, idx);
        const float sa = rng.next_f32(), sb = rng.next_f32();
        n_samples++;
        if (!kAdjoint) {
            DirectionSample ds = lens_sample_direction(L, L.ifs, src_p, sa, sb, A.origin_offset);
            if (!ds.valid) continue;
            n_valid++;
            FootprintD fp;
            if (!spot_footprint(F, ds.uv_x + (float) F.off_x, ds.uv_y + (float) F.off_y, fp, false)) continue;
            const float wv = ds.weight * A.scale;
            const float v0 = A.intensity[0] * wv, v1 = A.intensity[1] * wv, v2 = A.intensity[2] * wv, v3 = A.scale;
            for (int iy = 0; iy < fp.ny; ++iy) {
                const int y = fp.y0 + iy;
                for (int ix = 0; ix < fp.nx; ++ix) {
                    const int x = fp.x0 + ix;
                    const float w = fp.wx[ix] * fp.wy[iy];
                    if (w == 0.f) continue;
                    const int lx = x - A.win_x0, ly = y - A.win_y0;
                    float *dst = ((unsigned) lx < (unsigned) kWin && (unsigned) ly < (unsigned) kWin)
                                     ? win + (ly * kWin + lx) * 4 : A.accum + ((size_t) y * sw + x) * 4;
                    atomicAdd(dst, v0 * w); atomicAdd(dst + 1, v1 * w); atomicAdd(dst + 2, v2 * w); atomicAdd(dst + 3, v3 * w);
                }
            }
        } else {
            // ---- strict forward replay with a tape (identical masks to the forward kernel)
            using namespace strict;
            if (!(src_p.z >= L.near_clip && src_p.z <= L.far_clip)) continue;
            const LensIf &front = L.ifs[n_if - 1];
            float dx, dy; square_to_disk_concentric(sa, sb, dx, dy);
            V3 ap = v3(mul(dx, front.elem_radius), mul(dy, front.elem_radius), front.z);
            V3 dirv = sub(src_p, ap);
            float dist = norm(dirv), inv_dist = rcp(dist);
            V3 dir = mul(dirv, inv_dist);
            LensRay ray; ray.o = A.origin_offset > 0.f ? at(ap, dir, A.origin_offset) : src_p; ray.d = neg(dir);
            float tape[kMaxInterfaces][6];
            if (!lens_trace_from_world_taped(L.ifs, n_if, ray, tape)) continue;
            float t = div(-ray.o.z, ray.d.z);
            V3 pf = at(ray.o, ray.d, t);
            float sx = fmadd(L.f2s_m00, pf.x, L.f2s_m03), sy = fmadd(L.f2s_m11, pf.y, L.f2s_m13);
            if (!(sx >= 0.f && sy >= 0.f && sx <= 1.f && sy <= 1.f)) continue;
            n_valid++;
            const float ct_film = -ray.d.z;
            const float value = L.normalization / (ct_film * ct_film * ct_film * ct_film) * (L.rear_z * L.rear_z);
            const float wgt = value * inv_dist * inv_dist * dir.z;
            const float uvx = sx * L.res_x, uvy = sy * L.res_y;
            FootprintD fp;
            if (!spot_footprint(F, uvx + (float) F.off_x, uvy + (float) F.off_y, fp, true)) continue;
            // ---- transpose of the splat: adjoint of the sample weight and of the splat position
            const float wv = wgt * A.scale;
            const float v0 = A.intensity[0] * wv, v1 = A.intensity[1] * wv, v2 = A.intensity[2] * wv, v3 = A.scale;
            float S0 = 0.f, S1 = 0.f, S2 = 0.f, px_bar = 0.f, py_bar = 0.f;
            for (int iy = 0; iy < fp.ny; ++iy) {
                const int y = fp.y0 + iy;
                for (int ix = 0; ix < fp.nx; ++ix) {
                    const int x = fp.x0 + ix;
                    const float4 G = __ldg(reinterpret_cast<const float4 *>(A.grad_accum) + (size_t) y * sw + x);
                    const float w = fp.wx[ix] * fp.wy[iy];
                    S0 += G.x * w; S1 += G.y * w; S2 += G.z * w;
                    const float gv = G.x * v0 + G.y * v1 + G.z * v2 + G.w * v3;
                    px_bar += gv * fp.dwx[ix] * fp.wy[iy];
                    py_bar += gv * fp.wx[ix] * fp.dwy[iy];
                }
            }
            const float w_bar = A.scale * (A.intensity[0] * S0 + A.intensity[1] * S1 + A.intensity[2] * S2);
            // uv = s * res ; s = f2s * pf
            const float pfx_bar = px_bar * L.res_x * L.f2s_m00, pfy_bar = py_bar * L.res_y * L.f2s_m11;
            // pf = o + t d, t = -o.z / d.z
            const float3 o = f3(ray.o), d = f3(ray.d);
            AdjRay bar;
            bar.o = make_float3(pfx_bar, pfy_bar, 0.f);
            bar.d = make_float3(pfx_bar * t, pfy_bar * t, 0.f);
            const float t_bar = pfx_bar * d.x + pfy_bar * d.y;
            bar.o.z += -t_bar / d.z;
            bar.d.z += -t_bar * t / d.z;
            // wgt = value * dir.z / dist^2 ; value = norm * (-d.z)^-4 * rear_z^2
            const float value_bar = w_bar * inv_dist * inv_dist * dir.z;
            bar.d.z += value_bar * (-4.f * value / d.z);
            g_loc[0][0] += value_bar * 2.f * value / L.rear_z;            // rear_z = ifs[0].z
            float dirz_bar = w_bar * value * inv_dist * inv_dist;
            float dist_bar = w_bar * value * dir.z * (-2.f) * inv_dist * inv_dist * inv_dist;
            // ---- reverse sweep rear -> front
            for (int k = 0; k < n_if; ++k) {
                float g[kGradRow];
#pragma unroll
                for (int q = 0; q < kGradRow; ++q) g[q] = 0.f;
                // the hit point on interface k is the origin of the ray that entered interface k - 1 (or the final ray)
                const float3 hp = k > 0 ? make_float3(tape[k - 1][0], tape[k - 1][1], tape[k - 1][2]) : o;
                lens_interact_adj(L.ifs[k], make_float3(tape[k][0], tape[k][1], tape[k][2]),
                                  make_float3(tape[k][3], tape[k][4], tape[k][5]), hp, bar, g);
                g_loc[k][0] += g[G_Z]; g_loc[k][1] += g[G_R]; g_loc[k][2] += g[G_K];
#pragma unroll
                for (int q = 0; q < 7; ++q) g_loc[k][3 + q] += g[G_A0 + q];
                g_loc[k][10] += g[G_IORL]; g_loc[k][11] += g[G_IORR];
            }
            // ---- the ray's start: o0 = source (fixed), d0 = -dir, dir = (src - ap) / dist, ap.z = front.z
            // (with an origin offset: o0 = ap + offset * dir, the same line through ap)
            float3 dir_bar = make_float3(-bar.d.x, -bar.d.y, -bar.d.z + dirz_bar);
            float apz_bar = 0.f;
            if (A.origin_offset > 0.f) { dir_bar += bar.o * A.origin_offset; apz_bar = bar.o.z; }
            const float3 dirf = f3(dir);
            float3 v_bar = (dir_bar - dirf * dot(dirf, dir_bar)) * inv_dist + dirf * dist_bar;
            g_loc[n_if - 1][0] += apz_bar - v_bar.z;
        }
    }

    if (!kAdjoint) {
        __syncthreads();
        for (int i = tid; i < kWin * kWin; i += kSpotThreads) {
            const int ly = i / kWin, lx = i - ly * kWin, x = A.win_x0 + lx, y = A.win_y0 + ly;
            if ((unsigned) x >= (unsigned) sw || (unsigned) y >= (unsigned) sh) continue;
            const float4 v = reinterpret_cast<const float4 *>(win)[i];
            float *dst = A.accum + ((size_t) y * sw + x) * 4;
            if (v.x != 0.f) atomicAdd(dst, v.x);
            if (v.y != 0.f) atomicAdd(dst + 1, v.y);
            if (v.z != 0.f) atomicAdd(dst + 2, v.z);
            if (v.w != 0.f) atomicAdd(dst + 3, v.w);
        }
    } else {
        for (int i = 0; i < n_if; ++i)
            for (int k = 0; k < kSlots; ++k) {
                float v = g_loc[i][k];
                for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
                if (lane == 0 && v != 0.f) atomicAdd(&s_grad[i * kSlots + k], v);
            }
        __syncthreads();
        for (int i = tid; i < n_if * kSlots; i += kSpotThreads) {
            const float v = s_grad[i];
            if (v == 0.f) continue;
            const int r = i / kSlots, k = i - r * kSlots;
            const int col = k == 0 ? G_Z : (k == 1 ? G_R : (k == 2 ? G_K : (k < 10 ? G_A0 + (k - 3) : (k == 10 ? G_IORL : G_IORR))));
            atomicAdd(A.grad_rows + r * kGradRow + col, v);
        }
    }
    if (A.stats) {
        for (int o = 16; o > 0; o >>= 1) { n_samples += __shfl_down_sync(0xffffffffu, n_samples, o); n_valid += __shfl_down_sync(0xffffffffu, n_valid, o); }
        if (lane == 0) { atomicAdd(A.stats, n_samples); atomicAdd(A.stats + 1, n_valid); }
    }
}

static int spot_launch(lrt_sensor *s, const lrt_film_desc *film, const lrt_spot_params *sp, float *accum, const float *grad_accum,
                       float *grad_rows, unsigned long long *stats, cudaStream_t stream) {
    SpotArgs A; memset(&A, 0, sizeof A);
    lrt_film_desc fd = *film; fd.coalesce = 0; fd.normalize = 1;      // integrator.cpp:706-717
    int rc = film_params_from_desc(&fd, A.film); if (rc) return rc;
    LRT_REQUIRE(film->channels == 4, "spot images use the R,G,B,W storage of hdrfilm (channels = 4)");
    LRT_REQUIRE(A.film.kind != FILTER_BOX, "the spot adjoint needs a differentiable reconstruction filter (tent / gaussian)");
    LRT_REQUIRE(s->host.kind == 0, "spot rendering needs a realisticlens sensor");
    LRT_REQUIRE(sp->spp > 0 && sp->shard_count >= 1 && sp->shard_index >= 0 && sp->shard_index < sp->shard_count, "bad spot parameters");
    const double n_total = (double) film->size_x * film->size_y * sp->spp;
    LRT_REQUIRE(n_total < 4294967296.0, "W*H*spp exceeds 2^32: render in passes");
    rc = s->upload(stream); if (rc) return rc;
    A.lens = s->dev; A.seed = sp->seed; A.spp = sp->spp;
    for (int k = 0; k < 3; ++k) { A.src[k] = sp->src_p[k]; A.intensity[k] = sp->intensity[k]; }
    A.origin_offset = sp->origin_offset;
    const long long nt = (long long) n_total;
    A.n_my = (nt - sp->shard_index + sp->shard_count - 1) / sp->shard_count;
    A.shard_index = sp->shard_index; A.shard_count = sp->shard_count;
    A.scale = 1.f / (float) sp->spp;
    // window centre: the chief ray (centre of the front element), traced on the host with the same code
    DirectionSample c = lens_sample_direction(s->host, s->host.ifs, strict::v3(sp->src_p[0], sp->src_p[1], sp->src_p[2]), .5f, .5f, sp->origin_offset);
    float cx = c.valid ? c.uv_x : .5f * film->size_x, cy = c.valid ? c.uv_y : .5f * film->size_y;
    A.win_x0 = (int) floorf(cx) + film->border - kWin / 2; A.win_y0 = (int) floorf(cy) + film->border - kWin / 2;
    A.accum = accum; A.grad_accum = grad_accum; A.grad_rows = grad_rows; A.stats = stats;
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long want = (A.n_my + kSpotThreads - 1) / kSpotThreads;
    // every thread should see several samples so that the per-CTA window / gradient flush amortises
    long long grid = std::min<long long>((long long) sms * 8, std::max<long long>(1, want / 8));
    if (grad_rows) k_spot<true><<<(int) grid, kSpotThreads, 0, stream>>>(A);
    else k_spot<false><<<(int) grid, kSpotThreads, 0, stream>>>(A);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

__global__ void __launch_bounds__(256) k_sample_direction(const LensSystem *__restrict__ Lg, int64_t n, const float *__restrict__ p3,
                                                           const float2 *__restrict__ s2, float *__restrict__ out11) {
    __shared__ LensSystem L;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(Lg);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        DirectionSample ds = lens_sample_direction(L, L.ifs, strict::v3(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2]), s2[i].x, s2[i].y);
        float *o = out11 + 11 * i;
        o[0] = ds.uv_x; o[1] = ds.uv_y; o[2] = ds.p.x; o[3] = ds.p.y; o[4] = ds.p.z; o[5] = ds.d.x; o[6] = ds.d.y; o[7] = ds.d.z;
        o[8] = ds.dist; o[9] = ds.pdf; o[10] = ds.weight;
    }
}

}  // namespace lrt

using namespace lrt;
extern "C" {

int lrt_spot_render(const lrt_sensor *s, const lrt_film_desc *film, const lrt_spot_params *sp, float *accum_dev, uint64_t *stats_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && film && sp && accum_dev, "null argument");
    return spot_launch(const_cast<lrt_sensor *>(s), film, sp, accum_dev, nullptr, nullptr, (unsigned long long *) stats_dev, (cudaStream_t) stream);
}

int lrt_spot_render_backward(const lrt_sensor *s, const lrt_film_desc *film, const lrt_spot_params *sp, const float *grad_accum_dev,
                             float *grad_rows16_dev, uint64_t *stats_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && film && sp && grad_accum_dev && grad_rows16_dev, "null argument");
    return spot_launch(const_cast<lrt_sensor *>(s), film, sp, nullptr, grad_accum_dev, grad_rows16_dev, (unsigned long long *) stats_dev, (cudaStream_t) stream);
}

int lrt_sensor_sample_direction(const lrt_sensor *s_, int64_t n, const float *p3_dev, const float *s2_dev, float *out11_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s_ && n >= 0 && (n == 0 || (p3_dev && s2_dev && out11_dev)) && s_->host.kind == 0, "bad argument");
    if (n == 0) return LRT_OK;
    lrt_sensor *s = const_cast<lrt_sensor *>(s_);
    int rc = s->upload((cudaStream_t) stream); if (rc) return rc;
    int64_t g = (n + 255) / 256; if (g > 148 * 16) g = 148 * 16;
    k_sample_direction<<<(int) g, 256, 0, (cudaStream_t) stream>>>(s->dev, n, p3_dev, (const float2 *) s2_dev, out11_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_sensor_sample_direction_host(const lrt_sensor *s, int64_t n, const float *p3, const float *s2, float *out11) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && p3 && s2 && out11 && n >= 0, "bad argument");
    if (n == 0) return LRT_OK;
    float *d = nullptr;
    LRT_CUDA(cudaMalloc(&d, (size_t) n * 16 * sizeof(float)));
    cudaMemcpy(d, p3, (size_t) n * 12, cudaMemcpyHostToDevice);
    cudaMemcpy(d + 3 * n, s2, (size_t) n * 8, cudaMemcpyHostToDevice);
    int rc = lrt_sensor_sample_direction(s, n, d, d + 3 * n, d + 5 * n, nullptr);
    if (rc == LRT_OK) { cudaError_t e = cudaMemcpy(out11, d + 5 * n, (size_t) n * 44, cudaMemcpyDeviceToHost); if (e != cudaSuccess) rc = cuda_fail(e, "copy back"); }
    cudaFree(d);
    return rc;
}

// Host-side vector-Jacobian product of one lens trace (chief-ray sensitivities of the design front end; same
// LRT_HD code as the adjoint kernel): ray od6 is traced (from_world: front -> rear, else rear -> front), bar6
// is the adjoint of the outgoing ray; writes the adjoint of the incoming ray and ADDS d/d(rows16) to grad_rows16.
int lrt_sensor_trace_vjp_host(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6, const float *bar6,
                              float *out_bar6, float *grad_rows16, int32_t *active) {
    if (s) { int rc_ = const_cast<lrt_sensor *>(s)->sync_host(); if (rc_) return rc_; }
    LRT_REQUIRE(s && od6 && bar6 && out_bar6 && grad_rows16 && n >= 0 && s->host.kind == 0, "bad argument");
    const LensSystem &L = s->host;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = od6 + 6 * i;
        LensRay ray; ray.o = strict::v3(r[0], r[1], r[2]); ray.d = strict::v3(r[3], r[4], r[5]);
        float tape[kMaxInterfaces][6], hitp[kMaxInterfaces][3]; bool ok = true;
        for (int q = 0; q < L.n && ok; ++q) {
            int k = from_world ? L.n - 1 - q : q;
            tape[k][0] = ray.o.x; tape[k][1] = ray.o.y; tape[k][2] = ray.o.z; tape[k][3] = ray.d.x; tape[k][4] = ray.d.y; tape[k][5] = ray.d.z;
            ok = lens_interact(L.ifs[k], ray);
            hitp[k][0] = ray.o.x; hitp[k][1] = ray.o.y; hitp[k][2] = ray.o.z;
        }
        if (active) active[i] = ok ? 1 : 0;
        float *ob = out_bar6 + 6 * i;
        for (int k = 0; k < 6; ++k) ob[k] = 0.f;
        if (!ok) continue;
        AdjRay bar; bar.o = make_float3(bar6[6 * i], bar6[6 * i + 1], bar6[6 * i + 2]); bar.d = make_float3(bar6[6 * i + 3], bar6[6 * i + 4], bar6[6 * i + 5]);
        for (int q = L.n - 1; q >= 0; --q) {
            int k = from_world ? L.n - 1 - q : q;
            float g[kGradRow]; for (int c = 0; c < kGradRow; ++c) g[c] = 0.f;
            lens_interact_adj(L.ifs[k], make_float3(tape[k][0], tape[k][1], tape[k][2]), make_float3(tape[k][3], tape[k][4], tape[k][5]),
                              make_float3(hitp[k][0], hitp[k][1], hitp[k][2]), bar, g);
            for (int c = 0; c < kGradRow; ++c) grad_rows16[16 * k + c] += g[c];
        }
        ob[0] = bar.o.x; ob[1] = bar.o.y; ob[2] = bar.o.z; ob[3] = bar.d.x; ob[4] = bar.d.y; ob[5] = bar.d.z;
    }
    return LRT_OK;
}

}  // extern "C"
