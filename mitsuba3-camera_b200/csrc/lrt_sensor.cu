// lensrt — sensors: host-side lens construction (prescription -> interface table, thick-lens autofocus)
// and the unit-level sample_ray / sampler kernels behind include/lensrt.h.
// Host arithmetic that decides interface positions is float32 and strict (this TU is compiled with
// -Xcompiler -ffp-contract=off) because every later ray/lens comparison depends on those z's
// (realisticlens.cpp:1568-1651, :2107-2171).
#include <cstddef>
#include "lrt_internal.h"
#include "lrt_rng.cuh"
#include <cstring>
#include <cstdio>
#include <cmath>

namespace lrt {

static thread_local std::string g_error;
std::atomic<uint64_t> g_launches{0};

void set_error(const std::string &msg) { g_error = msg; }
int cuda_fail(cudaError_t e, const char *what) {
    g_error = std::string(what) + ": " + cudaGetErrorString(e);
    cudaGetLastError();
    return LRT_ECUDA;
}
bool have_device() {
    static int n = -1;
    if (n < 0) { int c = 0; if (cudaGetDeviceCount(&c) != cudaSuccess) { c = 0; cudaGetLastError(); } n = c; }
    return n > 0;
}

// ------------------------------------------------------------------ interface records
static float powi(float x, int y) {   // dr::pow(x, int): square-and-multiply (drjit/math.h:977-992)
    unsigned n = (unsigned) (y < 0 ? -y : y); float r = 1.f;
    while (n) { if (n & 1) r *= x; x *= x; n >>= 1; }
    return y >= 0 ? r : 1.f / r;
}

static float material_ior(const float *m, float wavelength) {   // realisticlens.cpp:226-240
    float w2 = wavelength * wavelength;
    if (m[0] == 0.f) return m[1] + m[2] / w2;
    float ior = 1.f;
    ior += m[1] * w2 / (w2 - m[4]);
    ior += m[2] * w2 / (w2 - m[5]);
    ior += m[3] * w2 / (w2 - m[6]);
    return sqrtf(ior);
}

// fill every derived constant of a record from its defining parameters (host and device: lrt_sensor_set_interfaces_dev)
LRT_HD static void finish_record(LensIf &f) {
    f.elem_r2 = f.elem_radius * f.elem_radius;
    f.eta = f.ior_right / f.ior_left;
    f.rcp_eta = 1.f / f.eta;
    f.center_z = f.z + f.curv_radius;
    f.curv_r2 = f.curv_radius * f.curv_radius;
    f.convex = f.curv_radius > 0.f ? 1 : 0;
    f.plane_param = -1.0f * f.z;
    f.rcp_R = 1.f / f.elem_radius;
    f.rcp_R2 = 1.f / (f.elem_radius * f.elem_radius);
    f.c2 = f.c * f.c;
    f.one_p_K = 1.f + f.K;
    f.k1 = f.one_p_K * f.c2;
    f.lim = f.K > -1.0f ? f.elem_radius * (1.f / (1.0f + f.K)) : inf_f();
    for (int i = 0; i < 7; ++i) f.gi[i] = (2.f * i + 4.f) * f.ai[i];
}

LRT_HD static LensIf blank(int kind, float z, float elem_radius, float il, float ir) {
    LensIf f; memset(&f, 0, sizeof f);
    f.kind = kind; f.z = z; f.elem_radius = elem_radius; f.ior_left = il; f.ior_right = ir;
    return f;
}

static void offset_along_axis(LensIf &f, float delta) { f.z += delta; finish_record(f); }

// 16-float row <-> record (layout documented in lensrt.h)
LRT_HD static LensIf record_from_row(const float *r) {
    int kind = (int) r[0];
    LensIf f = blank(kind, r[1], r[2], r[13], r[14]);
    if (kind == IF_SPHERE) f.curv_radius = r[3];
    if (kind == IF_ASPHERE) { f.c = r[3]; f.K = r[4]; f.r_scale = r[5]; for (int k = 0; k < 7; ++k) f.ai[k] = r[6 + k]; }
    if (kind == IF_APERTURE) f.ior_right = f.ior_left;
    finish_record(f);
    return f;
}
static void row_from_record(const LensIf &f, float *r) {
    for (int k = 0; k < 16; ++k) r[k] = 0.f;
    r[0] = (float) f.kind; r[1] = f.z; r[2] = f.elem_radius;
    if (f.kind == IF_SPHERE) r[3] = f.curv_radius;
    if (f.kind == IF_ASPHERE) { r[3] = f.c; r[4] = f.K; r[5] = f.r_scale; for (int k = 0; k < 7; ++k) r[6 + k] = f.ai[k]; }
    r[13] = f.ior_left; r[14] = f.ior_right;
}

// ------------------------------------------------------------------ film <-> sample transforms
// Film::get_physical_size (film.h:193-198) + film_to_crop_transform (sensor.h:305-342) and its inverse
// (realisticlens.cpp:2190-2207). The inverse is assembled from the inverses of the four factors, the way
// Transform::inverse() composes them, so the entries round like the reference's.
static void setup_film(lrt_sensor &s) {
    const lrt_film_geom &g = s.film;
    float diag = 0.001f * g.diagonal_mm;
    float aspect = (float) g.height / (float) g.width;
    float x = sqrtf(diag * diag / (1.f + aspect * aspect));
    float y = aspect * x;
    float px = x / 2, py = y / 2;
    float relx = (float) g.crop_w / (float) g.width, rely = (float) g.crop_h / (float) g.height;
    float offx = (float) g.crop_x / (float) g.width, offy = (float) g.crop_y / (float) g.height;
    float ax = 1.f / (1.f / relx), ay = 1.f / (1.f / rely);
    float cx = 1.f / (0.5f / px), cy = 1.f / (0.5f / py);
    float *m = s.sample_to_film;
    for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.f : 0.f;
    m[0] = ax * cx; m[5] = ay * cy;
    m[3] = offx * cx + (-px); m[7] = offy * cy + (-py);
    float *q = s.film_to_sample;
    for (int i = 0; i < 16; ++i) q[i] = (i % 5 == 0) ? 1.f : 0.f;
    float fx = (1.f / relx) * (0.5f / px), fy = (1.f / rely) * (0.5f / py);
    q[0] = fx; q[5] = fy;
    q[3] = fx * px + (1.f / relx) * (-offx); q[7] = fy * py + (1.f / rely) * (-offy);
    float x0 = m[3], y0 = m[7], x1 = m[0] + m[3], y1 = m[5] + m[7];
    s.normalization = 1.f / ((fmaxf(x0, x1) - fminf(x0, x1)) * (fmaxf(y0, y1) - fminf(y0, y1)));
    s.host.s2f_m00 = m[0]; s.host.s2f_m03 = m[3]; s.host.s2f_m11 = m[5]; s.host.s2f_m13 = m[7];
    s.host.persp_norm = s.normalization;
    s.host.f2s_m00 = q[0]; s.host.f2s_m03 = q[3]; s.host.f2s_m11 = q[5]; s.host.f2s_m13 = q[7];
    s.host.normalization = s.normalization; s.host.res_x = (float) g.crop_w; s.host.res_y = (float) g.crop_h;
    s.host.film_diagonal = sqrtf(px * px + py * py);   // m_film_diagonal = norm(get_physical_size()) (realisticlens.cpp:838)
}

// ------------------------------------------------------------------ thick-lens autofocus (:2107-2171)
static void cardinal_points(const LensRay &start, const LensRay &end, float &z_p, float &z_f) {
    float t_focus = -end.o.x / end.d.x;
    z_f = end.d.z * t_focus + end.o.z;
    float t_plane = (start.o.x - end.o.x) / end.d.x;
    z_p = end.d.z * t_plane + end.o.z;
}
static bool focus_thick_lens(const lrt_sensor &s, float half_x, float focus_distance, float &delta) {
    const LensSystem &L = s.host;
    float r = 0.001f * half_x;
    LensRay obj; obj.o = strict::v3(r, 0.f, L.ifs[L.n - 1].z + 1.f); obj.d = strict::v3(0.f, 0.f, -1.f);
    LensRay obj_end = obj;
    if (!lens_trace_from_world(L.ifs, L.n, obj_end)) return false;
    float obj_plane, obj_focus; cardinal_points(obj, obj_end, obj_plane, obj_focus);
    float p_img = obj_plane, f_img = obj_plane - obj_focus;
    LensRay img; img.o = strict::v3(r, 0.f, L.ifs[0].z - 1.f); img.d = strict::v3(0.f, 0.f, 1.f);
    LensRay img_end = img;
    if (!lens_trace_from_film(L.ifs, L.n, img_end)) return false;
    float img_plane, img_focus; cardinal_points(img, img_end, img_plane, img_focus);
    float p_obj = img_plane;
    float tmp = fmaxf(focus_distance, 4.01f * f_img) - p_obj;
    delta = 0.5f * (p_img - tmp + sqrtf((p_img + tmp) * (p_img + tmp) - 4 * f_img * (p_img + tmp)));
    return true;
}

static int finish_sensor(lrt_sensor *s, bool fine_focus, float object_distance) {
    LensSystem &L = s->host;
    if (fine_focus) {
        float aspect = (float) s->film.height / (float) s->film.width;
        float diag = 0.001f * s->film.diagonal_mm;
        float half_x = sqrtf(diag * diag / (1.f + aspect * aspect)) / 2;
        float delta;
        if (!focus_thick_lens(*s, half_x, object_distance, delta)) {
            set_error("compute_thick_lens_approximation: paraxial ray was not transmitted through lens");
            return LRT_ELENS;
        }
        s->focus_delta = delta;
        for (int i = 0; i < L.n; ++i) offset_along_axis(L.ifs[i], std::isnan(delta) ? 0.0f : -delta);
    }
    L.rear_z = L.ifs[0].z; L.rear_radius = L.ifs[0].elem_radius;
    s->dev_dirty = true;
    return LRT_OK;
}

static void set_common(lrt_sensor *s, const lrt_film_geom *film, const float *tw, float near_clip, float far_clip) {
    memset(&s->host, 0, sizeof s->host);
    s->film = *film;
    for (int i = 0; i < 12; ++i) s->host.to_world[i] = tw[i];
    s->host.near_clip = near_clip; s->host.far_clip = far_clip;
    s->host.pupil_lut = &s->host.pupil_bounds[0][0];
    setup_film(*s);
}

}  // namespace lrt

// the device table was written by lrt_sensor_set_interfaces_dev: bring the host copy up to date before it is read or re-uploaded
int lrt_sensor::sync_host() {
    if (!host_stale) return LRT_OK;
    LRT_CUDA(cudaDeviceSynchronize());
    LRT_CUDA(cudaMemcpy(host.ifs, reinterpret_cast<const char *>(dev) + offsetof(lrt::LensSystem, ifs), sizeof host.ifs, cudaMemcpyDeviceToHost));
    host.rear_z = host.ifs[0].z; host.rear_radius = host.ifs[0].elem_radius;
    host_stale = false;
    return LRT_OK;
}

int lrt_sensor::upload(cudaStream_t stream) {
    if (host_stale && dev_dirty) { int rc = sync_host(); if (rc) return rc; }
    if (!dev) LRT_CUDA(cudaMalloc(&dev, sizeof(lrt::LensSystem)));
    if (!staging) {
        LRT_CUDA(cudaMallocHost(&staging, sizeof(lrt::LensSystem)));
        LRT_CUDA(cudaEventCreateWithFlags(&staged, cudaEventDisableTiming));
    }
    host.pupil_lut = &host.pupil_bounds[0][0];
    if (dev_dirty) {
        LRT_CUDA(cudaEventSynchronize(staged));   // the previous upload has consumed the staging copy (no-op if none)
        *staging = host;   // the device copy's LUT pointer names the device copy's own array
        staging->pupil_lut = reinterpret_cast<const float *>(reinterpret_cast<const char *>(dev) + offsetof(lrt::LensSystem, pupil_bounds));
        LRT_CUDA(cudaMemcpyAsync(dev, staging, sizeof(lrt::LensSystem), cudaMemcpyHostToDevice, stream));
        LRT_CUDA(cudaEventRecord(staged, stream));
        dev_dirty = false;
    }
    return LRT_OK;
}

// ================================================================================ kernels
namespace lrt {

__global__ void __launch_bounds__(256) k_sample_ray(const LensSystem *__restrict__ Lg, int64_t n,
                                                     const float2 *__restrict__ pos, const float2 *__restrict__ ap,
                                                     float *__restrict__ out10) {
    __shared__ LensSystem L;
    {   // stage the lens table into shared memory (16-byte words)
        const uint4 *src = reinterpret_cast<const uint4 *>(Lg);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        float2 p = pos[i], q = ap[i];
        CameraRay r = sensor_sample_ray(L, L.ifs, p.x, p.y, q.x, q.y);
        float *o = out10 + 10 * i;
        o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
        o[6] = r.maxt; o[7] = o[8] = o[9] = r.weight;
    }
}

// the spectral variants' sample_ray / lens trace: one wavelength (nm) per ray, IORs from the media's dispersion formulas
__global__ void __launch_bounds__(256) k_sample_ray_wl(const LensSystem *__restrict__ Lg, const float *__restrict__ mat_lr, int64_t n,
                                                        const float2 *__restrict__ pos, const float2 *__restrict__ ap,
                                                        const float *__restrict__ wl, float *__restrict__ out10) {
    __shared__ LensSystem L;
    __shared__ float M[kMaxInterfaces * 14];
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(Lg);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
        for (int i = threadIdx.x; i < Lg->n * 14; i += blockDim.x) M[i] = mat_lr[i];
    }
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        float2 p = pos[i], q = ap[i];
        CameraRay r = lens_sample_ray_wl(L, L.ifs, M, wl[i], p.x, p.y, q.x, q.y);
        float *o = out10 + 10 * i;
        o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
        o[6] = r.maxt; o[7] = o[8] = o[9] = r.weight;
    }
}
__global__ void __launch_bounds__(256) k_trace_wl(const LensSystem *__restrict__ Lg, const float *__restrict__ mat_lr, int from_world, int64_t n,
                                                   const float *__restrict__ od6, const float *__restrict__ wl, float *__restrict__ out7) {
    __shared__ LensSystem L;
    __shared__ float M[kMaxInterfaces * 14];
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(Lg);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
        for (int i = threadIdx.x; i < Lg->n * 14; i += blockDim.x) M[i] = mat_lr[i];
    }
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float *r = od6 + 6 * i;
        LensRay ray; ray.o = strict::v3(r[0], r[1], r[2]); ray.d = strict::v3(r[3], r[4], r[5]);
        const bool a = lens_trace_wl(L.ifs, L.n, ray, M, wl[i], from_world != 0);
        float *o = out7 + 7 * i;
        o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.d.x; o[4] = ray.d.y; o[5] = ray.d.z; o[6] = a ? 1.f : 0.f;
    }
}

__global__ void __launch_bounds__(256) k_sampler_draw(uint32_t seed, uint32_t idx0, int64_t n, int ndraw, float *__restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        Pcg32 rng; rng.seed_sample(seed, idx0 + (uint32_t) i);
        for (int k = 0; k < ndraw; ++k) out[i * ndraw + k] = rng.next_f32();
    }
}

// bound_exit_pupil (realisticlens.cpp:1685-1767) for all 64 film segments at once: blockIdx.y = segment, the 2^20 Halton
// rays of a segment are dealt over blockIdx.x. The reference walks them in order and skips rays whose rear-plane point is
// already inside the box found so far; the box is the min / max over TRANSMITTED points either way, so the order does not
// matter. out[seg] = { min.x, min.y, max.x, max.y as order-preserving uints, transmitted count }.
LRT_D uint32_t ordered_bits(float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
// kWl: the spectral variants' LUT (realisticlens.cpp:1718-1727): every ray of a segment carries the wavelength
// lerp(380, 700, RadicalInverse(2, i)) nm and refracts with the media's IORs at that wavelength
template <bool kWl>
__global__ void __launch_bounds__(256) k_exit_pupil(const LensSystem *__restrict__ Lg, const float *__restrict__ mat_lr, uint32_t *__restrict__ out) {
    __shared__ LensSystem L;
    __shared__ float M[kWl ? kMaxInterfaces * 14 : 1];
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(Lg);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
        if (kWl) for (int i = threadIdx.x; i < Lg->n * 14; i += blockDim.x) M[i] = mat_lr[i];
    }
    __syncthreads();
    using namespace strict;
    const int seg = blockIdx.y;
    const uint32_t n_rays = 1u << kPupilRaysLog2;
    const float r0 = div(mul((float) seg, L.film_diagonal), (float) kPupilSegments),
                r1 = div(mul((float) (seg + 1), L.film_diagonal), (float) kPupilSegments);
    const float R = mul(L.ifs[0].elem_radius, 1.5f), rear_z = L.ifs[0].z;
    uint32_t mnx = 0xffffffffu, mny = 0xffffffffu, mxx = 0u, mxy = 0u, cnt = 0u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_rays; i += gridDim.x * blockDim.x) {
        const float fx = lerp_drjit(r0, r1, div(add((float) i, .5f), (float) n_rays));
        const float px = lerp_drjit(-R, R, radical_inverse(0, i)), py = lerp_drjit(-R, R, radical_inverse(1, i));
        LensRay ray; ray.o = v3(fx, 0.f, 0.f); ray.d = normalize(sub(v3(px, py, rear_z), ray.o));
        const bool through = kWl ? lens_trace_wl(L.ifs, L.n, ray, M, lerp_drjit(380.f, 700.f, radical_inverse(2, i)), false)
                                 : lens_trace_from_film(L.ifs, L.n, ray);
        if (through) {
            mnx = min(mnx, ordered_bits(px)); mxx = max(mxx, ordered_bits(px));
            mny = min(mny, ordered_bits(py)); mxy = max(mxy, ordered_bits(py));
            cnt++;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, o));
        mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, o)); mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if ((threadIdx.x & 31) == 0 && cnt) {
        uint32_t *q = out + 5 * seg;
        atomicMin(q, mnx); atomicMin(q + 1, mny); atomicMax(q + 2, mxx); atomicMax(q + 3, mxy); atomicAdd(q + 4, cnt);
    }
}

static float from_ordered_bits(uint32_t u) { u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u; float f; memcpy(&f, &u, 4); return f; }

// compute_exit_pupil_bounds (realisticlens.cpp:1769-1809): fills host.pupil_bounds from the current interfaces
static int upload_materials(lrt_sensor *s, cudaStream_t stream) {
    LRT_REQUIRE(s->host.kind == 0 && (int) s->mat_lr.size() == 14 * s->host.n,
                "this sensor has no dispersion data: build it from a prescription or call lrt_sensor_set_materials");
    LRT_REQUIRE(!s->host.pupil_mode || s->pupil_spectral,
                "exit-pupil sampling with per-ray wavelengths needs the spectral LUT: enable it with lrt_sensor_set_exit_pupil_spectral");
    if (!s->d_mat_lr) {
        LRT_CUDA(cudaMalloc(&s->d_mat_lr, sizeof(float) * 14 * kMaxInterfaces));
        LRT_CUDA(cudaMemcpyAsync(s->d_mat_lr, s->mat_lr.data(), sizeof(float) * s->mat_lr.size(), cudaMemcpyHostToDevice, stream));
        LRT_CUDA(cudaStreamSynchronize(stream));        // mat_lr is pageable host memory
    }
    return LRT_OK;
}

static int build_exit_pupil(lrt_sensor *s, cudaStream_t stream, bool spectral = false) {
    LRT_REQUIRE(s->host.kind == 0 && s->host.n > 0, "exit-pupil sampling needs a lens sensor");
    if (spectral) { const int was_mode = s->host.pupil_mode; s->host.pupil_mode = 0; int rc_ = upload_materials(s, stream); s->host.pupil_mode = was_mode; if (rc_) return rc_; }
    const int was = s->host.pupil_mode;
    s->host.pupil_mode = 0; s->dev_dirty = true;      // the build itself traces plain film -> world rays
    int rc = s->upload(stream); if (rc) return rc;
    uint32_t init[kPupilSegments][5], *d = nullptr;
    for (auto &r : init) { r[0] = r[1] = 0xffffffffu; r[2] = r[3] = r[4] = 0u; }
    LRT_CUDA(cudaMalloc(&d, sizeof init));
    cudaError_t e = cudaMemcpyAsync(d, init, sizeof init, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) {
        if (spectral) k_exit_pupil<true><<<dim3(148, kPupilSegments), 256, 0, stream>>>(s->dev, s->d_mat_lr, d);
        else k_exit_pupil<false><<<dim3(148, kPupilSegments), 256, 0, stream>>>(s->dev, nullptr, d);
        LRT_LAUNCHED();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(init, d, sizeof init, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(d);
    if (e != cudaSuccess) { s->host.pupil_mode = was; return cuda_fail(e, "exit-pupil LUT"); }
    const float R = s->host.ifs[0].elem_radius * 1.5f;
    const float spacing = 4 * R * sqrtf(2.f / (float) (1u << kPupilRaysLog2));   // :1762 (its own TODO notes the extra factor)
    for (int i = 0; i < kPupilSegments; ++i) {
        float mn[2] = { -R, -R }, mx[2] = { R, R };                              // zero transmission: the whole sampled square
        if (init[i][4]) { mn[0] = from_ordered_bits(init[i][0]); mn[1] = from_ordered_bits(init[i][1]);
                          mx[0] = from_ordered_bits(init[i][2]); mx[1] = from_ordered_bits(init[i][3]); }
        float *b = s->host.pupil_bounds[i];
        b[0] = mn[0] - spacing; b[1] = mn[1] - spacing; b[2] = mx[0] + spacing; b[3] = mx[1] + spacing;
    }
    s->host.pupil_mode = was; s->dev_dirty = true;
    return LRT_OK;
}

static int grid_for(int64_t n, int block) {
    int64_t g = (n + block - 1) / block;
    int64_t cap = 148 * 16;
    return (int) (g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace lrt

using namespace lrt;

// ================================================================================ C ABI
extern "C" {

const char *lrt_last_error(void) { return g_error.c_str(); }
int lrt_version(void) { return 100; }
uint64_t lrt_launch_count(void) { return g_launches.load(); }
int lrt_device_count(void) { int c = 0; if (cudaGetDeviceCount(&c) != cudaSuccess) { cudaGetLastError(); return 0; } return c; }
int lrt_set_device(int device) { LRT_NEED_DEVICE(); LRT_CUDA(cudaSetDevice(device)); return LRT_OK; }

static int create_prescription(int32_t n, const int32_t *kinds, const float *thick, const float *curv,
                               const float *elemr, const float *kappa, const float *ai7, const float *mats7,
                               float scale, float stop_ratio, int32_t fine_focus, float object_distance,
                               const lrt_film_geom *film, const float *tw, float near_clip, float far_clip,
                               float nominal_um, lrt_sensor **out) {
    LRT_REQUIRE(n > 0 && n <= kMaxInterfaces, "lens must have 1..32 interfaces");
    LRT_REQUIRE(kinds && thick && curv && elemr && kappa && ai7 && mats7 && film && tw && out, "null argument");
    lrt_sensor *s = new lrt_sensor();
    set_common(s, film, tw, near_clip, far_clip);
    LensSystem &L = s->host;
    // build_lens_from_data (:1568-1651): walk the table rear -> front accumulating z
    float z_pos = 0.f;
    for (int i = n - 1; i >= 0; --i) {
        float elem_radius = elemr[i];
        z_pos += thick[i];
        float il = material_ior(mats7 + 7 * (i + 1), nominal_um), ir = material_ior(mats7 + 7 * i, nominal_um);
        // the two media of this interface, for the per-ray IOR of the dispersion mode (the stop sits inside its left medium)
        s->mat_lr.insert(s->mat_lr.end(), mats7 + 7 * (i + 1), mats7 + 7 * (i + 1) + 7);
        { const float *right = kinds[i] == IF_APERTURE ? mats7 + 7 * (i + 1) : mats7 + 7 * i; s->mat_lr.insert(s->mat_lr.end(), right, right + 7); }
        LensIf f;
        switch (kinds[i]) {
            case IF_APERTURE: f = blank(IF_APERTURE, 0.001f * scale * z_pos, 0.001f * scale * elem_radius * stop_ratio, il, il); break;
            case IF_PLANE: f = blank(IF_PLANE, 0.001f * scale * z_pos, 0.001f * scale * elem_radius, il, ir); break;
            case IF_ASPHERE: {
                f = blank(IF_ASPHERE, 0.001f * scale * z_pos, 0.001f * scale * elem_radius, il, ir);
                // AsphericalLens ctor (:557-572): coefficients to unitless radius r/R
                f.K = kappa[i];
                f.r_scale = 1.0f / 0.001f * f.elem_radius;
                f.c = f.r_scale * (1.f / (scale * curv[i]));
                for (int j = 0; j < 7; ++j) f.ai[j] = powi(f.r_scale, 2 * j + 3) * (ai7[7 * i + j] / powi(scale, 2 * j + 3));
                break;
            }
            default: f = blank(IF_SPHERE, 0.001f * scale * z_pos, 0.001f * scale * elem_radius, il, ir);
                     f.curv_radius = -0.001f * scale * curv[i];
        }
        finish_record(f);
        L.ifs[L.n++] = f;
    }
    int rc = finish_sensor(s, fine_focus != 0, object_distance);
    if (rc != LRT_OK) { delete s; return rc; }
    *out = s;
    return LRT_OK;
}

int lrt_sensor_create_prescription(int32_t n, const int32_t *kinds, const float *thick, const float *curv,
                                   const float *elemr, const float *kappa, const float *ai7, const float *mats7,
                                   float scale, float stop_ratio, int32_t fine_focus, float object_distance,
                                   const lrt_film_geom *film, const float *tw, float near_clip, float far_clip,
                                   lrt_sensor **out) {
    return create_prescription(n, kinds, thick, curv, elemr, kappa, ai7, mats7, scale, stop_ratio, fine_focus, object_distance, film, tw,
                               near_clip, far_clip, 0.5893f, out);              // compute_ior(Float(0.5893f)), :170-173
}
int lrt_sensor_create_prescription_spectral(int32_t n, const int32_t *kinds, const float *thick, const float *curv,
                                            const float *elemr, const float *kappa, const float *ai7, const float *mats7,
                                            float scale, float stop_ratio, int32_t fine_focus, float object_distance,
                                            const lrt_film_geom *film, const float *tw, float near_clip, float far_clip,
                                            lrt_sensor **out) {
    // the spectral variants focus with rays of Wavelength(589.3f) nm (:2122-2146), i.e. IORs at 0.001f * 589.3f um — one ulp
    // above the rgb variants' 0.5893f
    return create_prescription(n, kinds, thick, curv, elemr, kappa, ai7, mats7, scale, stop_ratio, fine_focus, object_distance, film, tw,
                               near_clip, far_clip, 0.001f * 589.3f, out);
}

int lrt_sensor_create_interfaces(int32_t n, const float *rows16, float film_z, int32_t fine_focus, float object_distance,
                                 const lrt_film_geom *film, const float *tw, float near_clip, float far_clip,
                                 lrt_sensor **out) {
    LRT_REQUIRE(n > 0 && n <= kMaxInterfaces, "lens must have 1..32 interfaces");
    LRT_REQUIRE(rows16 && film && tw && out, "null argument");
    lrt_sensor *s = new lrt_sensor();
    set_common(s, film, tw, near_clip, far_clip);
    for (int i = 0; i < n; ++i) s->host.ifs[i] = record_from_row(rows16 + 16 * i);
    s->host.n = n; s->host.film_z = film_z;
    int rc = finish_sensor(s, fine_focus != 0, object_distance);
    if (rc != LRT_OK) { delete s; return rc; }
    *out = s;
    return LRT_OK;
}

int lrt_sensor_create_perspective(float x_fov_deg, const lrt_film_geom *film, const float *tw, float near_clip,
                                  float far_clip, lrt_sensor **out) {
    LRT_REQUIRE(film && tw && out, "null argument");
    lrt_sensor *s = new lrt_sensor();
    set_common(s, film, tw, near_clip, far_clip);
    s->host.kind = 1; s->persp_fov = x_fov_deg;
    // inverse of perspective_projection (sensor.h:262-303): S(1/rel) T(-off) S(-.5,-.5a,1) T(-1,-1/a,0) P(fov)
    double W = film->width, H = film->height, aspect = W / H;
    double relx = film->crop_w / W, rely = film->crop_h / H, offx = film->crop_x / W, offy = film->crop_y / H;
    double recip = 1.0 / ((double) far_clip - (double) near_clip);
    double tanh_ = std::tan(0.5 * (double) x_fov_deg * 3.14159265358979323846 / 180.0);
    // sample (u,v,0) -> ndc: x_c = (u*relx + offx) / -0.5 + 1 ; y_c = (v*rely + offy)/(-0.5*aspect) + 1/aspect
    // P^-1 at depth-buffer value 0 (near plane): camera = (x_c*tan, y_c*tan, 1) * near
    double n_ = near_clip;
    float *m = s->host.persp_s2c;
    for (int i = 0; i < 16; ++i) m[i] = 0.f;
    m[0] = (float) (relx / -0.5 * tanh_ * n_); m[3] = (float) ((offx / -0.5 + 1.0) * tanh_ * n_);
    m[5] = (float) (rely / (-0.5 * aspect) * tanh_ * n_); m[7] = (float) ((offy / (-0.5 * aspect) + 1.0 / aspect) * tanh_ * n_);
    m[11] = (float) n_; m[15] = 1.f;
    (void) recip;
    // image-plane normalisation (perspective.cpp update_camera_transforms): rect at z = 1
    double x0 = (offx / -0.5 + 1.0) * tanh_, x1 = ((relx + offx) / -0.5 + 1.0) * tanh_;
    double y0 = (offy / (-0.5 * aspect) + 1.0 / aspect) * tanh_, y1 = ((rely + offy) / (-0.5 * aspect) + 1.0 / aspect) * tanh_;
    s->host.persp_norm = (float) (1.0 / (std::fabs(x1 - x0) * std::fabs(y1 - y0)));
    s->dev_dirty = true;
    *out = s;
    return LRT_OK;
}

void lrt_sensor_destroy(lrt_sensor *s) {
    if (!s) return;
    if (s->dev) cudaFree(s->dev);
    if (s->d_mat_lr) cudaFree(s->d_mat_lr);
    if (s->staging) cudaFreeHost(s->staging);
    if (s->staged) cudaEventDestroy(s->staged);
    delete s;
}
int lrt_sensor_num_interfaces(const lrt_sensor *s) { return s ? s->host.n : 0; }
int lrt_sensor_get_interfaces(const lrt_sensor *s, float *rows16) {
    LRT_REQUIRE(s && rows16, "null argument");
    { int rc = const_cast<lrt_sensor *>(s)->sync_host(); if (rc) return rc; }
    for (int i = 0; i < s->host.n; ++i) row_from_record(s->host.ifs[i], rows16 + 16 * i);
    return LRT_OK;
}
namespace lrt {
__global__ void k_set_interfaces(LensSystem *L, int n, const float *__restrict__ rows16) {
    const int i = threadIdx.x;
    if (i < n) L->ifs[i] = record_from_row(rows16 + 16 * i);
    if (i == 0) { L->n = n; L->rear_z = rows16[1]; L->rear_radius = rows16[2]; }
}
}
// parameters_changed() for optimisation loops that keep the interface table on the DEVICE (rows16_dev [n,16] float32, same
// layout as lrt_sensor_set_interfaces): one small kernel on `stream`, no host round trip. The host copy is refreshed lazily
// (lrt_sensor_get_interfaces, the *_host entry points). Not available while exit-pupil sampling is on (its LUT is rebuilt
// from the host copy).
int lrt_sensor_set_interfaces_dev(lrt_sensor *s, int32_t n, const float *rows16_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && rows16_dev && n > 0 && n <= kMaxInterfaces && n <= 32, "bad argument");
    LRT_REQUIRE(s->host.kind == 0 && !s->host.pupil_mode, "set_interfaces_dev needs a realisticlens sensor in rear-element sampling mode");
    int rc = s->upload((cudaStream_t) stream); if (rc) return rc;
    lrt::k_set_interfaces<<<1, 32, 0, (cudaStream_t) stream>>>(s->dev, n, rows16_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    s->host.n = n; s->host_stale = true;
    return LRT_OK;
}
int lrt_sensor_set_interfaces(lrt_sensor *s, int32_t n, const float *rows16) {
    LRT_REQUIRE(s && rows16 && n > 0 && n <= kMaxInterfaces, "bad argument");
    s->host_stale = false;
    for (int i = 0; i < n; ++i) s->host.ifs[i] = record_from_row(rows16 + 16 * i);
    s->host.n = n; s->host.rear_z = s->host.ifs[0].z; s->host.rear_radius = s->host.ifs[0].elem_radius;
    s->dev_dirty = true;
    if (s->host.pupil_mode) return build_exit_pupil(s, nullptr);   // the LUT belongs to the interfaces it was traced through
    return LRT_OK;
}
// sample_exit_pupil = true|false (realisticlens.cpp:913-921) and the shutter interval the pupil-mode weight is scaled by
// (Sensor::shutter_open_time(), sensor.h; 0 unless the scene sets shutter_open / shutter_close)
int lrt_sensor_set_exit_pupil(lrt_sensor *s, int32_t enable, float shutter_open_time, void *stream) {
    LRT_REQUIRE(s, "null argument");
    s->host.shutter_open_time = shutter_open_time;
    s->dev_dirty = true;
    if (!enable) { s->host.pupil_mode = 0; return LRT_OK; }
    LRT_NEED_DEVICE();
    int rc = build_exit_pupil(s, (cudaStream_t) stream); if (rc) return rc;
    s->host.pupil_mode = 1; s->pupil_spectral = false; s->dev_dirty = true;
    return LRT_OK;
}
int lrt_sensor_set_exit_pupil_spectral(lrt_sensor *s, int32_t enable, float shutter_open_time, void *stream) {
    LRT_REQUIRE(s, "null argument");
    s->host.shutter_open_time = shutter_open_time;
    s->dev_dirty = true;
    if (!enable) { s->host.pupil_mode = 0; s->pupil_spectral = false; return LRT_OK; }
    LRT_NEED_DEVICE();
    int rc = build_exit_pupil(s, (cudaStream_t) stream, true); if (rc) return rc;
    s->host.pupil_mode = 1; s->pupil_spectral = true; s->dev_dirty = true;
    return LRT_OK;
}
int lrt_sensor_get_exit_pupil(const lrt_sensor *s, float *bounds256) {
    LRT_REQUIRE(s && bounds256, "null argument");
    LRT_REQUIRE(s->host.pupil_mode, "exit-pupil sampling is not enabled on this sensor");
    memcpy(bounds256, s->host.pupil_bounds, sizeof s->host.pupil_bounds);
    return LRT_OK;
}
float lrt_radical_inverse(int32_t base_index, uint64_t index) {
    return (base_index < 0 || base_index > 2) ? -1.f : lrt::radical_inverse(base_index, index);
}
int lrt_sensor_set_to_world(lrt_sensor *s, const float *tw) {
    LRT_REQUIRE(s && tw, "null argument");
    for (int i = 0; i < 12; ++i) s->host.to_world[i] = tw[i];
    s->dev_dirty = true;
    return LRT_OK;
}
int lrt_sensor_get_info(const lrt_sensor *s, float *s2f16, float *focus_delta, float *film_z) {
    LRT_REQUIRE(s, "null argument");
    if (s2f16) memcpy(s2f16, s->sample_to_film, 64);
    if (focus_delta) *focus_delta = s->focus_delta;
    if (film_z) *film_z = s->host.film_z;
    return LRT_OK;
}

int lrt_sensor_sample_ray(const lrt_sensor *s_, int64_t n, const float *pos2, const float *ap2, float *out10, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s_ && n >= 0 && (n == 0 || (pos2 && ap2 && out10)), "bad argument");
    if (n == 0) return LRT_OK;
    lrt_sensor *s = const_cast<lrt_sensor *>(s_);
    int rc = s->upload((cudaStream_t) stream); if (rc) return rc;
    k_sample_ray<<<grid_for(n, 256), 256, 0, (cudaStream_t) stream>>>(s->dev, n, (const float2 *) pos2, (const float2 *) ap2, out10);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_sensor_set_materials(lrt_sensor *s, int32_t n, const float *mat_lr14) {
    LRT_REQUIRE(s && mat_lr14 && s->host.kind == 0 && n == s->host.n, "lrt_sensor_set_materials: one (left, right) pair of 7-float rows per interface");
    for (int i = 0; i < 2 * n; ++i)
        LRT_REQUIRE(mat_lr14[7 * i] == 0.f || mat_lr14[7 * i] == 1.f, "material row: first entry 0 (Cauchy) or 1 (Sellmeier)");
    s->mat_lr.assign(mat_lr14, mat_lr14 + 14 * (size_t) n);
    if (s->d_mat_lr) { LRT_CUDA(cudaDeviceSynchronize()); LRT_CUDA(cudaFree(s->d_mat_lr)); s->d_mat_lr = nullptr; }
    return LRT_OK;
}
int lrt_sensor_get_materials(const lrt_sensor *s, float *mat_lr14) {
    LRT_REQUIRE(s && mat_lr14 && (int) s->mat_lr.size() == 14 * s->host.n, "this sensor has no dispersion data");
    memcpy(mat_lr14, s->mat_lr.data(), sizeof(float) * s->mat_lr.size());
    return LRT_OK;
}

int lrt_sensor_sample_ray_wl(const lrt_sensor *s_, int64_t n, const float *pos2, const float *ap2, const float *wl_nm, float *out10, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s_ && n >= 0 && (n == 0 || (pos2 && ap2 && wl_nm && out10)), "bad argument");
    if (n == 0) return LRT_OK;
    lrt_sensor *s = const_cast<lrt_sensor *>(s_);
    int rc = s->upload((cudaStream_t) stream); if (rc) return rc;
    rc = upload_materials(s, (cudaStream_t) stream); if (rc) return rc;
    k_sample_ray_wl<<<grid_for(n, 256), 256, 0, (cudaStream_t) stream>>>(s->dev, s->d_mat_lr, n, (const float2 *) pos2, (const float2 *) ap2, wl_nm, out10);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}
int lrt_sensor_trace_wl(const lrt_sensor *s_, int32_t from_world, int64_t n, const float *od6, const float *wl_nm, float *out7, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s_ && n >= 0 && (n == 0 || (od6 && wl_nm && out7)), "bad argument");
    if (n == 0) return LRT_OK;
    lrt_sensor *s = const_cast<lrt_sensor *>(s_);
    int rc = s->upload((cudaStream_t) stream); if (rc) return rc;
    rc = upload_materials(s, (cudaStream_t) stream); if (rc) return rc;
    k_trace_wl<<<grid_for(n, 256), 256, 0, (cudaStream_t) stream>>>(s->dev, s->d_mat_lr, from_world, n, od6, wl_nm, out7);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_sensor_sample_ray_host(const lrt_sensor *s, int64_t n, const float *pos2, const float *ap2, float *out10) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && n >= 0 && (n == 0 || (pos2 && ap2 && out10)), "bad argument");
    if (n == 0) return LRT_OK;
    float *d = nullptr;
    LRT_CUDA(cudaMalloc(&d, (size_t) n * 14 * sizeof(float)));
    cudaMemcpy(d, pos2, (size_t) n * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(d + 2 * n, ap2, (size_t) n * 8, cudaMemcpyHostToDevice);
    int rc = lrt_sensor_sample_ray(s, n, d, d + 2 * n, d + 4 * n, nullptr);
    if (rc == LRT_OK) { cudaError_t e = cudaMemcpy(out10, d + 4 * n, (size_t) n * 40, cudaMemcpyDeviceToHost); if (e != cudaSuccess) rc = cuda_fail(e, "copy back"); }
    cudaFree(d);
    return rc;
}

// host-side trace through the interface table (same LRT_HD code as the kernels; used by the autofocus and
// exposed for the lens-design front end's chief-ray queries: design_problem.py:415-417 uses prb_basic.trace)
int lrt_sensor_trace_host(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6, float *out7) {
    if (s) { int rc_ = const_cast<lrt_sensor *>(s)->sync_host(); if (rc_) return rc_; }
    LRT_REQUIRE(s && od6 && out7 && n >= 0 && s->host.kind == 0, "bad argument");
    for (int64_t i = 0; i < n; ++i) {
        const float *r = od6 + 6 * i;
        LensRay ray; ray.o = strict::v3(r[0], r[1], r[2]); ray.d = strict::v3(r[3], r[4], r[5]);
        bool a = from_world ? lens_trace_from_world(s->host.ifs, s->host.n, ray) : lens_trace_from_film(s->host.ifs, s->host.n, ray);
        float *o = out7 + 7 * i;
        o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.d.x; o[4] = ray.d.y; o[5] = ray.d.z; o[6] = a ? 1.f : 0.f;
    }
    return LRT_OK;
}

// host-side twin of lrt_sensor_trace_wl (same LRT_HD code), like lrt_sensor_trace_host
int lrt_sensor_trace_wl_host(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6, const float *wl_nm, float *out7) {
    if (s) { int rc_ = const_cast<lrt_sensor *>(s)->sync_host(); if (rc_) return rc_; }
    LRT_REQUIRE(s && od6 && wl_nm && out7 && n >= 0 && s->host.kind == 0, "bad argument");
    LRT_REQUIRE((int) s->mat_lr.size() == 14 * s->host.n, "this sensor has no dispersion data: build it from a prescription or call lrt_sensor_set_materials");
    for (int64_t i = 0; i < n; ++i) {
        const float *r = od6 + 6 * i;
        LensRay ray; ray.o = strict::v3(r[0], r[1], r[2]); ray.d = strict::v3(r[3], r[4], r[5]);
        bool a = lens_trace_wl(s->host.ifs, s->host.n, ray, s->mat_lr.data(), wl_nm[i], from_world != 0);
        float *o = out7 + 7 * i;
        o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.d.x; o[4] = ray.d.y; o[5] = ray.d.z; o[6] = a ? 1.f : 0.f;
    }
    return LRT_OK;
}

int lrt_sampler_draw(uint32_t seed, uint32_t idx0, int64_t n, int32_t ndraw, float *out, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(out && n >= 0 && ndraw > 0, "bad argument");
    if (n == 0) return LRT_OK;
    k_sampler_draw<<<grid_for(n, 256), 256, 0, (cudaStream_t) stream>>>(seed, idx0, n, ndraw, out);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}
int lrt_sampler_draw_host(uint32_t seed, uint32_t idx0, int64_t n, int32_t ndraw, float *out) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(out && n >= 0 && ndraw > 0, "bad argument");
    if (n == 0) return LRT_OK;
    float *d = nullptr;
    LRT_CUDA(cudaMalloc(&d, (size_t) n * ndraw * sizeof(float)));
    int rc = lrt_sampler_draw(seed, idx0, n, ndraw, d, nullptr);
    if (rc == LRT_OK) { cudaError_t e = cudaMemcpy(out, d, (size_t) n * ndraw * 4, cudaMemcpyDeviceToHost); if (e != cudaSuccess) rc = cuda_fail(e, "copy back"); }
    cudaFree(d);
    return rc;
}

}  // extern "C"
