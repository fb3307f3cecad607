// oracle/_ref harness — TEST INFRASTRUCTURE ONLY (never linked into / imported by the product).
//
// Compiles the REFERENCE's own code for the lens step and the RNG, from the sources where they
// lie under /root/reference, as the scalar_rgb variant (Float=float, Spectrum=Color<float,3>):
//   * header-only pieces are #included directly (core/random.h TEA, drjit/random.h PCG32,
//     render/fresnel.h, core/warp.h, core/math.h, render/sensor.h film_to_crop_transform);
//   * the lens classes and the camera's lens methods live inside one .cpp
//     (src/sensors/realisticlens.cpp) whose class derives from ProjectiveCamera and so cannot be
//     linked without libmitsuba; oracle/Makefile therefore pulls LINE RANGES of that file into
//     oracle/_ref/*.inc (generated, git-ignored) and this file supplies only the glue around
//     them: a struct carrying the members those methods touch, a Properties stand-in, and the
//     flat C entry points that tests/ call through ctypes.
// Everything between the "#include ref_*.inc" lines is the reference's text, untouched except
// for one sed substitution documented in the Makefile (shutter_open_time()).
#include <mitsuba/render/sensor.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/core/bbox.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/core/random.h>
#include <mitsuba/core/math.h>
#include <mitsuba/render/fresnel.h>
#include <drjit/random.h>
#include <map>
#include <string>
#include <sstream>
#include <iostream>
#include <cstring>
#include <memory>
#include <vector>

// RadicalInverse (core/qmc.h) keeps its base table private and builds it in src/core/qmc.cpp, which cannot be linked
// here; its eval() is header-only. The glue below fills the table exactly as qmc.cpp:46-52 does, hence the access.
#define private public
#include <mitsuba/core/qmc.h>
#undef private

// the glue reads a few private fields of the lens classes for the parameter dump
#define private public
#define protected public

NAMESPACE_BEGIN(mitsuba)

// Throw()/Log() outside a class name this global (defined in src/core/class.cpp as nullptr)
const Class *m_class = nullptr;
// RadicalInverse::eval()'s Throw() names the class's own m_class (src/core/qmc.cpp, not linked here)
Class *RadicalInverse::m_class = nullptr;
// src/core/logger.cpp's detail::Throw formats and throws std::runtime_error; same here
namespace detail {
void (Throw)(LogLevel, const Class *, const char *, int, const std::string &msg) {
    throw std::runtime_error(msg);
}
}

#include "_ref/ref_classes.inc"   // realisticlens.cpp:147-821

/// Stand-in for mitsuba::Properties: typed get<T>(name, default) over string maps.
struct PropsStub {
    std::map<std::string, std::string> s;
    std::map<std::string, double> f;
    template <typename T> T get(const std::string &name, const T &def) const {
        if constexpr (std::is_same_v<T, std::string>) {
            auto it = s.find(name); return it == s.end() ? def : it->second;
        } else {
            auto it = f.find(name); return it == f.end() ? def : T(it->second);
        }
    }
};

template <typename Float, typename Spectrum>
struct RefLensCamera {
    MI_IMPORT_TYPES()

    // ---- members the extracted methods touch (names as in realisticlens.cpp:2422-2438) ----
    struct FilmStub {   // Film::size/crop_size/crop_offset/get_physical_size (render/film.h:193-198)
        ScalarVector2u m_size, m_crop_size; ScalarPoint2u m_crop_offset; float m_diagonal;
        const ScalarVector2u &size() const { return m_size; }
        const ScalarVector2u &crop_size() const { return m_crop_size; }
        const ScalarPoint2u &crop_offset() const { return m_crop_offset; }
        ScalarVector2f get_physical_size() const {
            float aspect = float(m_size.y()) / float(m_size.x());
            float x = std::sqrt(m_diagonal * m_diagonal / (1.f + aspect * aspect));
            float y = aspect * x;
            return ScalarVector2f(x / 2, y / 2);
        }
    } film_stub, *m_film = &film_stub;
    BoundingBox2f m_image_rect;
    struct WorldStub {
        Transform4f v;
        const Transform4f &value() const { return v; }
    } m_to_world;
    Transform4f m_film_to_sample, m_sample_to_film;
    std::vector<LensInterface<Float, Spectrum> *> m_interfaces;
    Float m_rear_element_z = 0, m_rear_element_radius = 0, m_lens_terminal_z = 0;
    Float m_film_z_position = 0, m_stopdown_ratio = 1, m_film_diagonal = 0, m_normalization = 0;
    Float m_near_clip = 1e-2f, m_far_clip = 1e4f, m_shutter_open_time = 0;
    ScalarVector2f m_resolution;
    bool m_sample_exit_pupil = false;
    Float m_ray_wavelength = 589.3f;   // harness only: the wavelength (nm) sample_ray() gives its ray in the spectral instantiation

    // the reference dumps every focus ray to stdout here (realisticlens.cpp:2028-2099); silenced
    void draw_ray_from_world(const Ray3f &) const {}
    void draw_ray_from_film(const Ray3f &) const {}
    // exit-pupil LUT (realisticlens.cpp:2430-2436): the bounds, their flat copies, the Halton generator
    std::vector<BoundingBox2f> m_exit_pupil_bounds;
    DynamicBuffer<Float> m_min_bounds_ptr, m_max_bounds_ptr;
    RadicalInverse *m_qmc_sampler = nullptr;

    #include "_ref/ref_builders.inc"     // :936-1683  build_*_lens, build_lens_from_data, trace_ray_from_film
    #include "_ref/ref_pupil.inc"        // :1685-1845 bound_exit_pupil, compute_exit_pupil_bounds, sample_exit_pupil
    #include "_ref/ref_rear.inc"         // :1870-1878 sample_rear_element
    #include "_ref/ref_trace_world.inc"  // :1960-1984 trace_ray_from_world
    #include "_ref/ref_transforms.inc"   // :2190-2207 update_camera_transforms
    #include "_ref/ref_focus.inc"        // :2107-2171 cardinal points, thick-lens approx, focus_thick_lens

    void init(const PropsStub &props) {
        m_film_z_position = 0.0f;
        #include "_ref/ref_ctor_body.inc"   // :844-912 lens selection, fine focus, rear/terminal z
    }

    std::pair<Ray3f, Spectrum> sample_ray(Float time, const Point2f &position_sample,
                                          const Point2f &aperture_sample, Mask active) const {
        Spectrum wav_weight(1.f);   // sample_wavelengths() in *_rgb: weight 1, no wavelengths
        Ray3f ray;
        ray.time = time;
        // spectral instantiation: realisticlens.cpp:2215-2221 draws the wavelengths (sample_wavelengths: the film's sensor
        // response, outside this path) and stores them in the ray; the test passes the wavelength in
        if constexpr (is_spectral_v<Spectrum>) ray.wavelengths = Wavelength(m_ray_wavelength);
        #include "_ref/ref_sample_ray_body.inc"   // :2223-2291
    }

    std::pair<DirectionSample3f, Spectrum>
    sample_direction(const Interaction3f &it, const Point2f &sample, Mask active) const {
        #include "_ref/ref_sample_dir_body.inc"   // :2304-2393
    }
};

NAMESPACE_END(mitsuba)

using namespace mitsuba;
using Cam = RefLensCamera<float, Color<float, 3>>;
using P2 = Point<float, 2>;
using P3 = Point<float, 3>;
using V3 = Vector<float, 3>;
using RayF = mitsuba::Ray<P3, Color<float, 3>>;

struct CoutSilencer {
    std::ostringstream sink; std::streambuf *old;
    CoutSilencer() { old = std::cout.rdbuf(sink.rdbuf()); }
    ~CoutSilencer() { std::cout.rdbuf(old); }
};

extern "C" {

// ---- RNG (core/random.h:76-90, drjit/random.h:43-137) ---------------------------------------
void ref_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) {
    auto [a, b] = sample_tea_32<uint32_t>(v0, v1, rounds);
    out[0] = a; out[1] = b;
}
float ref_tea_float32(uint32_t v0, uint32_t v1, int rounds) {
    return sample_tea_float32<uint32_t>(v0, v1, rounds);
}
void ref_pcg32_u32(uint64_t initstate, uint64_t initseq, int n, uint32_t *out) {
    dr::PCG32<uint32_t> rng(1, initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = rng.next_uint32();
}
void ref_pcg32_f32(uint64_t initstate, uint64_t initseq, int n, float *out) {
    dr::PCG32<uint32_t> rng(1, initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = rng.next_float32();
}

// ---- small maths (fresnel.h:35-84,311-315; warp.h:54-90; math.h:360-402) --------------------
void ref_fresnel(int n, const float *cos_i, const float *eta, float *out4) {
    for (int i = 0; i < n; ++i) {
        auto [r, ct, eit, eti] = fresnel<float>(cos_i[i], eta[i]);
        out4[4 * i] = r; out4[4 * i + 1] = ct; out4[4 * i + 2] = eit; out4[4 * i + 3] = eti;
    }
}
void ref_refract(int n, const float *wi, const float *m, const float *ct, const float *eti, float *out3) {
    for (int i = 0; i < n; ++i) {
        V3 r = refract(V3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]),
                       Normal<float, 3>(m[3 * i], m[3 * i + 1], m[3 * i + 2]), ct[i], eti[i]);
        out3[3 * i] = r.x(); out3[3 * i + 1] = r.y(); out3[3 * i + 2] = r.z();
    }
}
void ref_concentric_disk(int n, const float *uv, float *out2) {
    for (int i = 0; i < n; ++i) {
        P2 p = warp::square_to_uniform_disk_concentric(P2(uv[2 * i], uv[2 * i + 1]));
        out2[2 * i] = p.x(); out2[2 * i + 1] = p.y();
    }
}
void ref_cosine_hemisphere(int n, const float *uv, float *out3) {
    for (int i = 0; i < n; ++i) {
        V3 p = warp::square_to_cosine_hemisphere(P2(uv[2 * i], uv[2 * i + 1]));
        out3[3 * i] = p.x(); out3[3 * i + 1] = p.y(); out3[3 * i + 2] = p.z();
    }
}
void ref_solve_quadratic(int n, const float *abc, float *out3) {
    for (int i = 0; i < n; ++i) {
        auto [v, x0, x1] = math::solve_quadratic(abc[3 * i], abc[3 * i + 1], abc[3 * i + 2]);
        out3[3 * i] = v ? 1.f : 0.f; out3[3 * i + 1] = x0; out3[3 * i + 2] = x1;
    }
}

// ---- camera ----------------------------------------------------------------------------------
// film: width, height, crop (w,h,ox,oy), diagonal in mm (film.cpp:46-48: m_diagonal = 0.001*diag)
}   // extern "C"
template <typename Cam>
static void *cam_create(const char *lens_design, int width, int height, const int *crop4,
                        float diagonal_mm, const float *to_world16, float near_clip, float far_clip,
                        float shutter_open_time, int n_props, const char **prop_names,
                        const double *prop_values) {
    CoutSilencer quiet;
    Cam *c = new Cam();
    c->film_stub.m_size = { (uint32_t) width, (uint32_t) height };
    c->film_stub.m_crop_size = { (uint32_t) crop4[0], (uint32_t) crop4[1] };
    c->film_stub.m_crop_offset = { (uint32_t) crop4[2], (uint32_t) crop4[3] };
    c->film_stub.m_diagonal = 0.001f * diagonal_mm;
    c->m_resolution = { (float) crop4[0], (float) crop4[1] };
    dr::Matrix<float, 4> m;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m(i, j) = to_world16[4 * i + j];
    c->m_to_world.v = typename Cam::Transform4f(m);
    c->m_near_clip = near_clip; c->m_far_clip = far_clip; c->m_shutter_open_time = shutter_open_time;
    // ctor order of realisticlens.cpp:832-842
    c->update_camera_transforms();
    c->m_film_diagonal = dr::norm(c->m_film->get_physical_size());
    PropsStub props;
    props.s["lens_design"] = lens_design;
    for (int i = 0; i < n_props; ++i) props.f[prop_names[i]] = prop_values[i];
    try { c->init(props); } catch (const std::exception &e) {
        fprintf(stderr, "ref_cam_create: %s\n", e.what()); delete c; return nullptr;
    }
    return c;
}
// the same lens classes and camera methods instantiated as a SPECTRAL variant (Spectrum = mitsuba::Spectrum<float, 4>):
// DispersiveMaterial::compute_ior(ray) then evaluates the media at 0.001 * ray.wavelengths[0] (realisticlens.cpp:169-179)
using SCam = RefLensCamera<float, mitsuba::Spectrum<float, 4>>;
using SRayF = mitsuba::Ray<Point<float, 3>, mitsuba::Spectrum<float, 4>>;
extern "C" {
void *ref_cam_create(const char *lens_design, int width, int height, const int *crop4,
                     float diagonal_mm, const float *to_world16, float near_clip, float far_clip,
                     float shutter_open_time, int n_props, const char **prop_names,
                     const double *prop_values) {
    return cam_create<Cam>(lens_design, width, height, crop4, diagonal_mm, to_world16, near_clip, far_clip, shutter_open_time, n_props,
                           prop_names, prop_values);
}
void *ref_scam_create(const char *lens_design, int width, int height, const int *crop4,
                      float diagonal_mm, const float *to_world16, float near_clip, float far_clip,
                      float shutter_open_time, int n_props, const char **prop_names,
                      const double *prop_values) {
    return cam_create<SCam>(lens_design, width, height, crop4, diagonal_mm, to_world16, near_clip, far_clip, shutter_open_time, n_props,
                            prop_names, prop_values);
}
void ref_scam_destroy(void *h) { delete (SCam *) h; }
int ref_scam_num_interfaces(void *h) { return (int) ((SCam *) h)->m_interfaces.size(); }
float ref_scam_film_z(void *h) { return ((SCam *) h)->m_film_z_position; }
// per interface z, element radius and the two IORs at the spectral variants' nominal wavelength (Wavelength(589.3f))
void ref_scam_interface(void *h, int i, float *out4) {
    auto *it = ((SCam *) h)->m_interfaces[i];
    SRayF ray(Point<float, 3>(0.f), Vector<float, 3>(0.f, 0.f, 1.f), 0.f, mitsuba::Spectrum<float, 4>(589.3f));
    out4[0] = it->m_z_intercept; out4[1] = it->m_element_radius;
    out4[2] = it->m_left_material.compute_ior(ray); out4[3] = it->m_right_material.compute_ior(ray);
}
// Sensor::sample_ray of the spectral instantiation, one wavelength (nm) per sample: out = o[3] d[3] maxt weight(first channel)
void ref_scam_sample_ray(void *h, int n, const float *pos2, const float *ap2, const float *wl_nm, float *out10) {
    SCam *c = (SCam *) h;
    for (int i = 0; i < n; ++i) {
        c->m_ray_wavelength = wl_nm[i];
        auto [ray, w] = c->sample_ray(0.f, P2(pos2[2 * i], pos2[2 * i + 1]), P2(ap2[2 * i], ap2[2 * i + 1]), true);
        float *o = out10 + 10 * i;
        o[0] = ray.o.x(); o[1] = ray.o.y(); o[2] = ray.o.z();
        o[3] = ray.d.x(); o[4] = ray.d.y(); o[5] = ray.d.z();
        o[6] = ray.maxt; o[7] = w[0]; o[8] = w[1]; o[9] = w[2];
    }
}
// trace_ray_from_film / trace_ray_from_world with the ray's wavelength
void ref_scam_trace(void *h, int from_world, int n, const float *od6, const float *wl_nm, float *out7) {
    SCam *c = (SCam *) h;
    for (int i = 0; i < n; ++i) {
        const float *r = od6 + 6 * i;
        SRayF ray(P3(r[0], r[1], r[2]), V3(r[3], r[4], r[5]), 0.f, mitsuba::Spectrum<float, 4>(wl_nm[i]));
        auto [ro, act] = from_world ? c->trace_ray_from_world(ray) : c->trace_ray_from_film(ray);
        float *o = out7 + 7 * i;
        o[0] = ro.o.x(); o[1] = ro.o.y(); o[2] = ro.o.z();
        o[3] = ro.d.x(); o[4] = ro.d.y(); o[5] = ro.d.z(); o[6] = act ? 1.f : 0.f;
    }
}
void ref_cam_destroy(void *h) { delete (Cam *) h; }
int ref_cam_num_interfaces(void *h) { return (int) ((Cam *) h)->m_interfaces.size(); }
float ref_cam_film_z(void *h) { return ((Cam *) h)->m_film_z_position; }
void ref_cam_sample_to_film(void *h, float *out16) {
    auto &m = ((Cam *) h)->m_sample_to_film.matrix;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out16[4 * i + j] = m(i, j);
}
// per interface: kind (0 sphere,1 plane,2 aperture,3 asphere), z, element radius, curvature radius /
// asphere c, kappa, r_scale, 7 coeffs, left ior, right ior @ 0.5893 um  -> 16 floats
void ref_cam_interface(void *h, int i, float *out16) {
    using F = float; using S = Color<float, 3>;
    auto *it = ((Cam *) h)->m_interfaces[i];
    for (int k = 0; k < 16; ++k) out16[k] = 0.f;
    out16[1] = it->m_z_intercept; out16[2] = it->m_element_radius;
    if (auto *s = dynamic_cast<SpheroidLens<F, S> *>(it)) { out16[0] = 0; out16[3] = s->m_curvature_radius; }
    else if (auto *a = dynamic_cast<ApertureStop<F, S> *>(it)) { (void) a; out16[0] = 2; }
    else if (auto *p = dynamic_cast<PlanoLens<F, S> *>(it)) { (void) p; out16[0] = 1; }
    else if (auto *a = dynamic_cast<AsphericalLens<F, S> *>(it)) {
        out16[0] = 3; out16[3] = a->m_c; out16[4] = a->m_K; out16[5] = a->m_r_scale;
        for (size_t k = 0; k < a->m_ai.size() && k < 7; ++k) out16[6 + k] = a->m_ai[k];
    }
    out16[13] = it->m_left_material.compute_ior(0.5893f);
    out16[14] = it->m_right_material.compute_ior(0.5893f);
}
float ref_material_ior(int cauchy, const float *coef, float wavelength_um) {
    using M = DispersiveMaterial<float, Color<float, 3>>;
    if (cauchy) return M("m", coef[0], coef[1]).compute_ior(wavelength_um);
    return M("m", V3(coef[0], coef[1], coef[2]), V3(coef[3], coef[4], coef[5])).compute_ior(wavelength_um);
}
// Exit-pupil LUT (realisticlens.cpp:913-921): `m_qmc_sampler = new RadicalInverse()` then compute_exit_pupil_bounds().
// The generator object is assembled without its constructor (src/core/qmc.cpp is not linkable here): the three prime
// bases eval() can be asked for (2, 3, 5) are filled as qmc.cpp:46-52 does. Returns 64 x (min.x, min.y, max.x, max.y).
}   // extern "C"
template <typename Cam>
static void cam_compute_exit_pupil(void *h, float *out256) {
    CoutSilencer quiet;
    Cam *c = (Cam *) h;
    if (!c->m_qmc_sampler) {
        RadicalInverse *q = (RadicalInverse *) ::operator new(sizeof(RadicalInverse));
        memset((void *) q, 0, sizeof(RadicalInverse));
        const uint64_t primes[3] = { 2, 3, 5 };
        auto *base = new RadicalInverse::PrimeBase[3];
        for (int i = 0; i < 3; ++i) {
            base[i].value = (uint16_t) primes[i];
            base[i].recip = 1.f / (float) primes[i];
            base[i].divisor = dr::divisor<uint64_t>(primes[i]);
        }
        new (&q->m_base) std::unique_ptr<RadicalInverse::PrimeBase[]>(base);
        q->m_base_count = 3;
        c->m_qmc_sampler = q;
    }
    c->m_sample_exit_pupil = true;
    c->compute_exit_pupil_bounds();
    for (size_t i = 0; i < c->m_exit_pupil_bounds.size() && i < 64; ++i) {
        const auto &b = c->m_exit_pupil_bounds[i];
        out256[4 * i] = b.min.x(); out256[4 * i + 1] = b.min.y(); out256[4 * i + 2] = b.max.x(); out256[4 * i + 3] = b.max.y();
    }
}
extern "C" {
void ref_cam_compute_exit_pupil(void *h, float *out256) { cam_compute_exit_pupil<Cam>(h, out256); }
// the spectral instantiation's LUT: every ray of the build carries lerp(380, 700, RadicalInverse(2, i)) nm (:1718-1727)
void ref_scam_compute_exit_pupil(void *h, float *out256) { cam_compute_exit_pupil<SCam>(h, out256); }
void ref_scam_set_shutter_open_time(void *h, float t) { ((SCam *) h)->m_shutter_open_time = t; }
float ref_radical_inverse(int base_index, uint64_t index) {
    static RadicalInverse *q = nullptr;
    if (!q) {
        q = (RadicalInverse *) ::operator new(sizeof(RadicalInverse));
        memset((void *) q, 0, sizeof(RadicalInverse));
        const uint64_t primes[3] = { 2, 3, 5 };
        auto *base = new RadicalInverse::PrimeBase[3];
        for (int i = 0; i < 3; ++i) { base[i].value = (uint16_t) primes[i]; base[i].recip = 1.f / (float) primes[i]; base[i].divisor = dr::divisor<uint64_t>(primes[i]); }
        new (&q->m_base) std::unique_ptr<RadicalInverse::PrimeBase[]>(base);
        q->m_base_count = 3;
    }
    return q->eval<float>((size_t) base_index, index);
}
// Sensor::sample_ray (realisticlens.cpp:2209-2292): out = o[3] d[3] maxt weight[3] per sample
void ref_cam_sample_ray(void *h, int n, const float *pos2, const float *ap2, float *out10) {
    Cam *c = (Cam *) h;
    for (int i = 0; i < n; ++i) {
        auto [ray, w] = c->sample_ray(0.f, P2(pos2[2 * i], pos2[2 * i + 1]), P2(ap2[2 * i], ap2[2 * i + 1]), true);
        float *o = out10 + 10 * i;
        o[0] = ray.o.x(); o[1] = ray.o.y(); o[2] = ray.o.z();
        o[3] = ray.d.x(); o[4] = ray.d.y(); o[5] = ray.d.z();
        o[6] = ray.maxt; o[7] = w.x(); o[8] = w.y(); o[9] = w.z();
    }
}
// trace_ray_from_film / trace_ray_from_world (:1654-1683 / :1960-1984): in o[3] d[3], out o[3] d[3] active
void ref_cam_trace(void *h, int from_world, int n, const float *od6, float *out7) {
    Cam *c = (Cam *) h;
    for (int i = 0; i < n; ++i) {
        const float *r = od6 + 6 * i;
        RayF ray(P3(r[0], r[1], r[2]), V3(r[3], r[4], r[5]));
        auto [ro, act] = from_world ? c->trace_ray_from_world(ray) : c->trace_ray_from_film(ray);
        float *o = out7 + 7 * i;
        o[0] = ro.o.x(); o[1] = ro.o.y(); o[2] = ro.o.z();
        o[3] = ro.d.x(); o[4] = ro.d.y(); o[5] = ro.d.z(); o[6] = act ? 1.f : 0.f;
    }
}
// Sensor::sample_direction (:2301-2394): in ref point p[3], sample[2];
// out uv[2] p[3] d[3] dist pdf weight  -> 11 floats
void ref_cam_sample_direction(void *h, int n, const float *p3, const float *s2, float *out11) {
    Cam *c = (Cam *) h;
    for (int i = 0; i < n; ++i) {
        Cam::Interaction3f it = dr::zeros<Cam::Interaction3f>();
        it.p = P3(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2]); it.t = 0.f;
        auto [ds, w] = c->sample_direction(it, P2(s2[2 * i], s2[2 * i + 1]), true);
        float *o = out11 + 11 * i;
        o[0] = ds.uv.x(); o[1] = ds.uv.y(); o[2] = ds.p.x(); o[3] = ds.p.y(); o[4] = ds.p.z();
        o[5] = ds.d.x(); o[6] = ds.d.y(); o[7] = ds.d.z(); o[8] = ds.dist; o[9] = ds.pdf; o[10] = w.x();
    }
}

// the same for the spectral instantiation (which, like the rgb one, traces the connection at Wavelength(589.3f):
// realisticlens.cpp:2339-2344 "TODO: account for effect of dispersion on ray path")
void ref_scam_sample_direction(void *h, int n, const float *p3, const float *s2, float *out11) {
    SCam *c = (SCam *) h;
    for (int i = 0; i < n; ++i) {
        SCam::Interaction3f it = dr::zeros<SCam::Interaction3f>();
        it.p = P3(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2]); it.t = 0.f;
        auto [ds, w] = c->sample_direction(it, P2(s2[2 * i], s2[2 * i + 1]), true);
        float *o = out11 + 11 * i;
        o[0] = ds.uv.x(); o[1] = ds.uv.y(); o[2] = ds.p.x(); o[3] = ds.p.y(); o[4] = ds.p.z();
        o[5] = ds.d.x(); o[6] = ds.d.y(); o[7] = ds.d.z(); o[8] = ds.dist; o[9] = ds.pdf; o[10] = w[0];
    }
}

}  // extern "C"
