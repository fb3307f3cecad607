// lensrt — the spot kernels (see lrt_spot.cu for what they compute). This header is compiled TWICE: by lrt_spot.cu with the
// strict arithmetic of lrt_math.cuh (bit-exact masks: the flavour the golden-vector tests pin) and by lrt_spot_fast.cu with
// LRT_FAST_ARITH defined (fused multiply-adds, MUFU reciprocal / square root: same algorithm, last-ulp differences, about half
// the instructions — the flavour design loops use, lrt_spot_params.fast). kFlavour only keeps the two instantiations' symbols
// apart; the arithmetic switch is the macro.
#pragma once
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
    const uint32_t *seed_dev; uint32_t seed_stride, seed_xor;   // seed = (seed + *seed_dev * seed_stride) ^ seed_xor when seed_dev is set
    float src[3], intensity[3], origin_offset;
    long long n_my; int shard_index, shard_count;
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

template <bool kAdjoint, int kFlavour>
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
    // window centre = the chief ray (centre of the front element) traced through the table as it is ON THE DEVICE (design loops
    // update it there, lrt_sensor_set_interfaces_dev): every CTA derives the same origin
    __shared__ int s_win[2];
    if (!kAdjoint && tid == 0) {
        const DirectionSample c = lens_sample_direction(L, L.ifs, strict::v3(A.src[0], A.src[1], A.src[2]), .5f, .5f, A.origin_offset);
        const float cx = c.valid ? c.uv_x : .5f * (float) F.size_x, cy = c.valid ? c.uv_y : .5f * (float) F.size_y;
        s_win[0] = (int) floorf(cx) + F.border - kWin / 2; s_win[1] = (int) floorf(cy) + F.border - kWin / 2;
    }
    if (!kAdjoint) __syncthreads();
    const int win_x0 = kAdjoint ? 0 : s_win[0], win_y0 = kAdjoint ? 0 : s_win[1];
    const int sw = F.size_x + 2 * F.border, sh = F.size_y + 2 * F.border;
    const strict::V3 src_p = strict::v3(A.src[0], A.src[1], A.src[2]);
    const int n_if = L.n;
    // CUDA-graph friendly seeding: the iteration counter lives on the device, the launch parameters never change
    const uint32_t seed = A.seed_dev ? ((A.seed + __ldg(A.seed_dev) * A.seed_stride) ^ A.seed_xor) : A.seed;
    unsigned long long n_samples = 0, n_valid = 0;
    float g_loc[kAdjoint ? kMaxInterfaces : 1][kSlots];
    if (kAdjoint) for (int i = 0; i < n_if; ++i) for (int k = 0; k < kSlots; ++k) g_loc[i][k] = 0.f;

    for (long long i = (long long) blockIdx.x * kSpotThreads + tid; i < A.n_my; i += (long long) gridDim.x * kSpotThreads) {
        const uint32_t idx = (uint32_t) (A.shard_index + i * A.shard_count);
        Pcg32 rng; rng.seed_sample(seed, idx);
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
                    const int lx = x - win_x0, ly = y - win_y0;
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
            const int ly = i / kWin, lx = i - ly * kWin, x = win_x0 + lx, y = win_y0 + ly;
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


// launch of this translation unit's flavour
static inline cudaError_t LRT_SPOT_LAUNCH(const SpotArgs &A, bool adjoint, int grid, cudaStream_t stream) {
    if (adjoint) k_spot<true, LRT_SPOT_FLAVOUR><<<grid, kSpotThreads, 0, stream>>>(A);
    else k_spot<false, LRT_SPOT_FLAVOUR><<<grid, kSpotThreads, 0, stream>>>(A);
    return cudaGetLastError();
}

}  // namespace lrt
