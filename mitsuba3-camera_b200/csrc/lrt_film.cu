// lensrt — film entry points: standalone ImageBlock::put / HDRFilm::develop kernels (unit-level parity with
// src/render/tests/test_imageblock.py) and the FilmParams set-up shared with the render kernels.
#include "lrt_internal.h"
#include "lrt_film.cuh"
#include <cmath>

namespace lrt {

static float estrin_host(float x, const float *coeff, int n) {   // drjit/math.h:27-41
    float buf[16], nxt[16];
    for (int i = 0; i < n; ++i) buf[i] = coeff[i];
    while (n > 1) {
        int n_rec = (n - 1) / 2, n_fma = n / 2;
        for (int i = 0; i < n_fma; ++i) nxt[i] = x * buf[2 * i + 1] + buf[2 * i];
        if (n_rec == n_fma) nxt[n_rec] = buf[n - 1];
        n = n_rec + 1;
        for (int i = 0; i < n; ++i) buf[i] = nxt[i];
        x = x * x;
    }
    return buf[0];
}

int film_params_from_desc(const lrt_film_desc *d, FilmParams &F) {
    LRT_REQUIRE(d, "null film descriptor");
    LRT_REQUIRE(d->size_x > 0 && d->size_y > 0 && d->channels > 0 && d->channels <= 8 && d->border >= 0, "bad film size");
    LRT_REQUIRE(d->filter >= 0 && d->filter <= 2, "filter must be box(0) / tent(1) / gaussian(2)");
    // ImageBlock's border is the filter's border_size(): 0 for the box filter (imageblock.cpp:24-29)
    LRT_REQUIRE(d->filter != 0 || d->border == 0, "a box-filtered block has no border");
    memset(&F, 0, sizeof F);
    F.size_x = d->size_x; F.size_y = d->size_y; F.off_x = d->off_x; F.off_y = d->off_y; F.border = d->border;
    F.channels = d->channels; F.kind = d->filter; F.coalesce = d->coalesce; F.normalize = d->normalize;
    if (F.kind == FILTER_BOX) { F.radius = .5f; F.inv_radius = 2.f; }
    if (F.kind == FILTER_TENT) { F.radius = d->filter_param > 0 ? d->filter_param : 1.f; F.inv_radius = 1.f / F.radius; }
    if (F.kind == FILTER_GAUSSIAN) {   // gaussian.cpp:48-89
        float stddev = d->filter_param > 0 ? d->filter_param : .5f;
        F.radius = 4 * stddev;
        const double c[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                               -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
        double scale = 1;
        for (int i = 0; i < 10; ++i) { F.coeff[i] = (float) (c[i] * scale); scale /= (double) stddev * (double) stddev; }
        F.coeff[0] -= estrin_host(F.radius * F.radius, F.coeff, 10);
    }
    for (int i = 0; i < 31; ++i) F.disc[i] = filter_eval(F, (F.radius * (float) i) / 31.f);
    F.disc[31] = 0.f; F.disc_scale = 31.f / F.radius; F.discretized = 0;
    F.n_half = (int) std::ceil(F.radius - .5f); F.count = 2 * F.n_half + 1;
    F.count_max = (int) std::ceil(2.f * F.radius);
    LRT_REQUIRE(F.count <= kMaxTaps && F.count_max <= kMaxTaps, "filter radius too large (max 8 taps per axis)");
    return LRT_OK;
}

__global__ void __launch_bounds__(256) k_film_put(FilmParams F, int64_t n, const float2 *__restrict__ pos,
                                                   const float *__restrict__ values, float *__restrict__ accum) {
    int sx = F.size_x + 2 * F.border, sy = F.size_y + 2 * F.border;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        float2 p = pos[i];
        Footprint fp;
        if (!film_footprint(F, p.x, p.y, fp)) continue;
        const float *v = values + i * F.channels;
        for (int iy = 0; iy < fp.ny; ++iy) {
            unsigned y = (unsigned) (fp.y0 + iy);
            if (y >= (unsigned) sy) continue;
            for (int ix = 0; ix < fp.nx; ++ix) {
                unsigned x = (unsigned) (fp.x0 + ix);
                if (x >= (unsigned) sx) continue;
                float w = fp.wy[iy] * fp.wx[ix];
                float *dst = accum + ((size_t) y * sx + x) * F.channels;
                for (int k = 0; k < F.channels; ++k) atomicAdd(dst + k, v[k] * w);
            }
        }
    }
}

// HDRFilm::develop, pixel_format rgb (hdrfilm.cpp:390-392): RGB / (W == 0 ? 1 : W)
__global__ void __launch_bounds__(256) k_develop_rgb(int64_t npix, const float4 *__restrict__ storage, float *__restrict__ rgb) {
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < npix; i += (int64_t) gridDim.x * blockDim.x) {
        float4 s = storage[i];
        float d = s.w == 0.f ? 1.f : s.w;
        rgb[3 * i] = s.x / d; rgb[3 * i + 1] = s.y / d; rgb[3 * i + 2] = s.z / d;
    }
}

}  // namespace lrt

using namespace lrt;
extern "C" {

int lrt_film_put(const lrt_film_desc *desc, int64_t n, const float *pos2_dev, const float *values_dev, float *accum_dev, void *stream) {
    LRT_NEED_DEVICE();
    FilmParams F; int rc = film_params_from_desc(desc, F); if (rc) return rc;
    LRT_REQUIRE(accum_dev && n >= 0 && (n == 0 || (pos2_dev && values_dev)), "bad argument");
    if (n == 0) return LRT_OK;
    int64_t g = (n + 255) / 256; if (g > 148 * 16) g = 148 * 16;
    k_film_put<<<(int) g, 256, 0, (cudaStream_t) stream>>>(F, n, (const float2 *) pos2_dev, values_dev, accum_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_film_develop_rgb(int64_t npix, const float *storage4_dev, float *rgb_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(storage4_dev && rgb_dev && npix >= 0, "bad argument");
    if (npix == 0) return LRT_OK;
    int64_t g = (npix + 255) / 256; if (g > 148 * 16) g = 148 * 16;
    k_develop_rgb<<<(int) g, 256, 0, (cudaStream_t) stream>>>(npix, (const float4 *) storage4_dev, rgb_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_filter_eval_host(int32_t filter, float param, int32_t n, const float *x, float *out, float *radius) {
    lrt_film_desc d; memset(&d, 0, sizeof d);
    d.size_x = d.size_y = 1; d.channels = 1; d.filter = filter; d.filter_param = param;
    FilmParams F; int rc = film_params_from_desc(&d, F); if (rc) return rc;
    for (int i = 0; i < n; ++i) out[i] = filter_eval(F, x[i]);
    if (radius) *radius = F.radius;
    return LRT_OK;
}
}
