// lensrt — micro-kernels that MEASURE the ceilings the render kernels are compared with ("FP32 peak is not in
// MEASURED_PEAKS.json -> measure with an FMA micro-kernel on the box"). Nothing on this path is HBM- or tensor-bound: the
// relevant denominators are the FP32 / issue rate of the SMs, the L1 and L2 sector rates and the shared-memory atomic rate.
// lrt_measure_peaks() runs each for a few milliseconds (CUDA events on the caller's stream) and returns
//   out[0] FP32 TFLOP/s (FFMA = 2 flop, 8 independent chains per thread, 148 x 4 x 256 threads)
//   out[1] warp instructions issued per second, G/s (the same kernel: one FFMA per issue slot)
//   out[2] L1 hit bandwidth, GB/s (128-bit loads over a 16 KB window per CTA)
//   out[3] L2 hit bandwidth, GB/s (128-bit ld.global.cg over a 32 MB window: larger than every L1, inside the 126 MB L2)
//   out[4] shared-memory float atomicAdd rate, G atomics/s (conflict-free addresses; a CAS loop on this architecture)
//   out[5] SM clock during the run, MHz (clock64 / event time)
#include "lrt_internal.h"

namespace lrt {

__global__ void __launch_bounds__(256) k_peak_fma(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 12345.678f) out[0] = s;   // never true: keeps the chains alive
}

__global__ void __launch_bounds__(256) k_peak_l1(const uint4 *__restrict__ buf, uint4 *out, int iters) {
    // every CTA walks its own 16 KB window (1024 x uint4): after the first pass all loads hit L1
    const uint4 *win = buf + (size_t) (blockIdx.x % 64) * 1024;
    uint4 acc = make_uint4(0, 0, 0, 0);
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 v = __ldg(win + ((threadIdx.x + 256 * k + i * 37) & 1023));
            acc.x ^= v.x; acc.y += v.y; acc.z ^= v.z; acc.w += v.w;
        }
    }
    if (acc.x == 0x12345678u && acc.y == 1u) out[0] = acc;
}

__global__ void __launch_bounds__(256) k_peak_l2(const uint4 *buf, uint4 *out, int iters, unsigned mask) {
    uint4 acc = make_uint4(0, 0, 0, 0);
    unsigned idx = (blockIdx.x * 256u + threadIdx.x) & mask;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 v = __ldcg(buf + idx);
            acc.x ^= v.x; acc.y += v.y; acc.z ^= v.z; acc.w += v.w;
            idx = (idx + 148u * 8u * 256u) & mask;
        }
    }
    if (acc.x == 0x12345678u && acc.y == 1u) out[0] = acc;
}

__global__ void __launch_bounds__(256) k_peak_smem_atomic(float *out, int iters) {
    __shared__ float s[256 * 4];
    for (int i = threadIdx.x; i < 1024; i += 256) s[i] = 0.f;
    __syncthreads();
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 4; ++k) atomicAdd(&s[threadIdx.x + 256 * ((k + i) & 3)], 1.f);
    }
    __syncthreads();
    if (s[threadIdx.x] == -1.f) out[0] = s[threadIdx.x];
}

__global__ void k_peak_clock(long long *out, int iters) {
    const long long t0 = clock64();
    float x = threadIdx.x;
#pragma unroll 1
    for (int i = 0; i < iters; ++i) x = fmaf(x, 1.0001f, 0.5f);
    out[0] = clock64() - t0; if (x == 1.2345f) out[1] = 0;
}

}  // namespace lrt

using namespace lrt;

extern "C" int lrt_measure_peaks(double *out6, void *stream_) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(out6, "null argument");
    cudaStream_t stream = (cudaStream_t) stream_;
    int dev = 0, sms = 148; LRT_CUDA(cudaGetDevice(&dev)); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t l2_bytes = 32u << 20;
    uint4 *buf = nullptr; float *sink = nullptr;
    LRT_CUDA(cudaMalloc(&buf, l2_bytes)); LRT_CUDA(cudaMalloc(&sink, 256));
    LRT_CUDA(cudaMemsetAsync(buf, 1, l2_bytes, stream));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto timed = [&](auto launch) -> double {   // best of 3, milliseconds
        double best = 1e30;
        for (int r = 0; r < 4; ++r) {
            cudaEventRecord(e0, stream); launch(); cudaEventRecord(e1, stream); cudaEventSynchronize(e1);
            float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < best) best = ms;
        }
        return best;
    };
    const int grid = sms * 8;
    {
        const int iters = 2048;
        const double ms = timed([&] { k_peak_fma<<<grid, 256, 0, stream>>>(sink, iters, 1.0001f, 0.5f); });
        const double ffma = (double) grid * 256 * iters * 16 * 8;
        out6[0] = 2.0 * ffma / (ms * 1e-3) / 1e12;
        out6[1] = ffma / 32.0 / (ms * 1e-3) / 1e9;
    }
    {
        const int iters = 4096;
        const double ms = timed([&] { k_peak_l1<<<grid, 256, 0, stream>>>(buf, (uint4 *) sink, iters); });
        out6[2] = (double) grid * 256 * iters * 4 * 16 / (ms * 1e-3) / 1e9;
    }
    {
        const int iters = 256;
        const double ms = timed([&] { k_peak_l2<<<grid, 256, 0, stream>>>(buf, (uint4 *) sink, iters, (unsigned) (l2_bytes / 16 - 1)); });
        out6[3] = (double) grid * 256 * iters * 4 * 16 / (ms * 1e-3) / 1e9;
    }
    {
        const int iters = 1024;
        const double ms = timed([&] { k_peak_smem_atomic<<<grid, 256, 0, stream>>>(sink, iters); });
        out6[4] = (double) grid * 256 * iters * 4 / (ms * 1e-3) / 1e9;
    }
    {
        long long *d = nullptr, h[2] = { 0, 0 };
        cudaMalloc(&d, 16);
        const double ms = timed([&] { k_peak_clock<<<1, 32, 0, stream>>>(d, 1 << 20); });
        cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost); cudaFree(d);
        out6[5] = (double) h[0] / (ms * 1e-3) / 1e6;
    }
    for (int k = 0; k < 5; ++k) LRT_LAUNCHED();
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf); cudaFree(sink);
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}
