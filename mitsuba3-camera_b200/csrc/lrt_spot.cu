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
#define LRT_SPOT_FLAVOUR 0
#define LRT_SPOT_LAUNCH launch_spot_flavour_strict
#include "lrt_spot_kernel.cuh"

namespace lrt {
cudaError_t launch_spot_fast(const SpotArgs &A, bool adjoint, int grid, cudaStream_t stream);   // lrt_spot_fast.cu

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
    A.seed_dev = sp->seed_dev; A.seed_stride = sp->seed_stride; A.seed_xor = sp->seed_xor;
    for (int k = 0; k < 3; ++k) { A.src[k] = sp->src_p[k]; A.intensity[k] = sp->intensity[k]; }
    A.origin_offset = sp->origin_offset;
    const long long nt = (long long) n_total;
    A.n_my = (nt - sp->shard_index + sp->shard_count - 1) / sp->shard_count;
    A.shard_index = sp->shard_index; A.shard_count = sp->shard_count;
    A.scale = 1.f / (float) sp->spp;
    A.accum = accum; A.grad_accum = grad_accum; A.grad_rows = grad_rows; A.stats = stats;
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long want = (A.n_my + kSpotThreads - 1) / kSpotThreads;
    // every thread should see several samples so that the per-CTA window / gradient flush amortises
    long long grid = std::min<long long>((long long) sms * 8, std::max<long long>(1, want / 8));
    LRT_CUDA(sp->fast ? launch_spot_fast(A, grad_rows != nullptr, (int) grid, stream) : launch_spot_flavour_strict(A, grad_rows != nullptr, (int) grid, stream));
    LRT_LAUNCHED();
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
