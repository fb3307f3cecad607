// lensrt — an entry point that allocates NOTHING: render + develop with the film accumulator in a fixed device array.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
//
// lrt_render_host (lrt_scene.cu) allocates its R,G,B,W accumulator and its RGB image with cudaMalloc on every call. Here the
// accumulator — the scratch of the render kernel — lives in the file-scope array below (4 MiB: at most 2^18 film pixels, e.g.
// 512 x 512), which every GPU that loads the library holds until exit; the caller passes the scene, the sensor, the film, the
// parameters and the DEVICE buffer the image goes to. Every call shares the array and the array keeps the last call's
// contents: the part a call uses is cleared on the caller's stream first, and two calls that overlap on the GPU (different
// streams without an ordering between them, or different host threads) RACE on it.
//
// In host code `g_noalloc_accum` is a host-side stand-in for the symbol, not a device address: the address the kernels get is
// looked up once per device, with that device current, by cudaGetSymbolAddress.
#include "lrt_scene.h"
#include <mutex>

namespace lrt {

constexpr int64_t kNoAllocMaxPixels = 1ll << 18;
static __device__ __align__(256) float g_noalloc_accum[kNoAllocMaxPixels * 4];

// device address of the scratch array on the CURRENT device (looked up on first use, then cached per device)
static int noalloc_scratch(float **out) {
    static std::mutex mu;
    static float *ptr[64] = {};
    int dev = 0;
    LRT_CUDA(cudaGetDevice(&dev));
    LRT_REQUIRE(dev >= 0 && dev < 64, "lrt_render_rgb_noalloc: device ordinal out of range (64 devices at most)");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        float *p = nullptr;
        LRT_CUDA(cudaGetSymbolAddress((void **) &p, g_noalloc_accum));
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return LRT_OK;
}

}  // namespace lrt

using namespace lrt;
extern "C" {

int64_t lrt_render_rgb_noalloc_max_pixels(void) { return kNoAllocMaxPixels; }

int lrt_render_rgb_noalloc(const lrt_scene *sc, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                           float *rgb_out_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && sensor && film && rp && rgb_out_dev, "null argument");
    LRT_REQUIRE(film->border == 0 && film->channels == 4, "lrt_render_rgb_noalloc needs a borderless R,G,B,W film");
    LRT_REQUIRE(film->size_x > 0 && film->size_y > 0, "empty film");
    const int64_t npix = (int64_t) film->size_x * film->size_y;
    if (npix > kNoAllocMaxPixels) {
        set_error("lrt_render_rgb_noalloc: the film has " + std::to_string(npix) + " pixels, the fixed scratch array holds at most " +
                  std::to_string(kNoAllocMaxPixels) + " (use lrt_render with a caller-owned accumulator for larger films)");
        return LRT_EINVAL;
    }
    LRT_REQUIRE(rp->rng_mode == 0, "lrt_render_rgb_noalloc: only the default sampler convention (rng_mode 0) runs without allocating");
    // the lens table's device mirror and its pinned staging copy are created by the sensor's FIRST upload; that is set-up work
    LRT_REQUIRE(sensor->dev && sensor->staging,
                "lrt_render_rgb_noalloc: the sensor has not been uploaded yet (its first use allocates its device table): "
                "call lrt_render or lrt_sensor_sample_ray once at set-up");
    float *accum = nullptr;
    int rc = noalloc_scratch(&accum); if (rc) return rc;
    LRT_CUDA(cudaMemsetAsync(accum, 0, (size_t) npix * 4 * sizeof(float), (cudaStream_t) stream));   // the last call's film
    rc = lrt_render(sc, sensor, film, rp, accum, stream); if (rc) return rc;
    return lrt_film_develop_rgb(npix, accum, rgb_out_dev, stream);
}

}  // extern "C"
