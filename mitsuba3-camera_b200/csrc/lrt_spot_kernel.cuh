// lensrt — the spot kernels (see lrt_spot.cu for what they compute). This header is compiled TWICE: by lrt_spot.cu with the
// strict arithmetic of lrt_math.cuh (bit-exact masks: the flavour the golden-vector tests pin) and by lrt_spot_fast.cu with
// LRT_FAST_ARITH defined (fused multiply-adds, MUFU reciprocal / square root: same algorithm, last-ulp differences, about half
// the instructions — the flavour design loops use, lrt_spot_params.fast). kFlavour only keeps the two instantiations' symbols
// apart; the arithmetic switch is the macro.
#pragma once
#include "lrt_internal.h"
#include "lrt_rng.cuh"
#include "lrt_lens_adj.cuh"
#include "lrt_spot_adj.cuh"

namespace lrt {

constexpr int kSpotThreads = 128;
constexpr int kWin = 32;            // shared window: kWin x kWin pixels x 4 channels = 16 KB

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
            if (lens_connection_adjoint(L, F, src_p, sa, sb, A.origin_offset, A.intensity[0], A.intensity[1], A.intensity[2], A.scale, A.scale,
                                        A.grad_accum, g_loc)) n_valid++;
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
