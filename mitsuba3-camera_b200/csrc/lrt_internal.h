// lensrt — internal declarations shared by the translation units of liblensrt.so
#pragma once
#include "../../include/lensrt.h"
#include "lrt_lens.cuh"
#include "lrt_film.cuh"
#include <atomic>
#include <string>
#include <vector>

namespace lrt {

void set_error(const std::string &msg);
int cuda_fail(cudaError_t e, const char *what);
bool have_device();
extern std::atomic<uint64_t> g_launches;
int film_params_from_desc(const lrt_film_desc *d, FilmParams &F);

#define LRT_CUDA(call)                                                            \
    do {                                                                          \
        cudaError_t e_ = (call);                                                  \
        if (e_ != cudaSuccess) return lrt::cuda_fail(e_, #call);                  \
    } while (0)
#define LRT_REQUIRE(cond, msg)                                                    \
    do {                                                                          \
        if (!(cond)) { lrt::set_error(msg); return LRT_EINVAL; }                  \
    } while (0)
#define LRT_NEED_DEVICE()                                                         \
    do {                                                                          \
        if (!lrt::have_device()) {                                                \
            lrt::set_error("no CUDA device: liblensrt has no CPU fallback");      \
            return LRT_ENODEVICE;                                                 \
        }                                                                         \
    } while (0)
#define LRT_LAUNCHED() (lrt::g_launches.fetch_add(1, std::memory_order_relaxed))

}  // namespace lrt

struct lrt_sensor {
    lrt::LensSystem host;            // authoritative copy
    lrt::LensSystem *dev = nullptr;  // lazily uploaded mirror (per current device)
    bool dev_dirty = true;
    lrt_film_geom film;
    float focus_delta = 0.f;
    float sample_to_film[16];
    float film_to_sample[16];
    float normalization = 0.f;
    lrt::LensSystem *staging = nullptr;   // pinned copy the asynchronous upload reads from
    cudaEvent_t staged = nullptr;         // completion of the last upload (the staging copy is reused after it)
    // host -> dev if dirty, ordered on `stream` (renders launched earlier on that stream keep reading the old table)
    int upload(cudaStream_t stream = nullptr);
    bool host_stale = false;              // the device table is newer than `host.ifs` (lrt_sensor_set_interfaces_dev)
    // dispersion mode (spectral variants): per interface the Cauchy / Sellmeier rows of its two media (left 7, right 7), in the
    // order of host.ifs; empty when the sensor was built from ready-made interfaces and lrt_sensor_set_materials was not called
    std::vector<float> mat_lr; float *d_mat_lr = nullptr;
    float persp_fov = 0.f;                // perspective sensors: x fov in degrees (the particle tracer rebuilds camera -> sample from it)
    bool pupil_spectral = false;          // the exit-pupil LUT in use was built with per-ray wavelengths (lrt_sensor_set_exit_pupil_spectral)
    int sync_host();
};
