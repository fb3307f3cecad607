// lensrt — the realistic-lens step: sequential trace of one ray through spherical / planar / aperture /
// aspheric interfaces with Snell refraction and vignetting.
// Replaces src/sensors/realisticlens.cpp: LensInterface::compute_interaction (:269-297), SpheroidLens
// (:359-402), PlanoLens (:444-473), ApertureStop (:505-517), AsphericalLens (:574-819),
// trace_ray_from_film (:1654-1683), trace_ray_from_world (:1960-1984), sample_rear_element (:1870-1878)
// and RealisticLensCamera::sample_ray (:2209-2292); with fresnel()/refract() of
// include/mitsuba/render/fresnel.h:35-84,311-315, square_to_uniform_disk_concentric of
// include/mitsuba/core/warp.h:54-90 and Dr.Jit's sincos (ext/drjit/include/drjit/math.h:76-176).
//
// Design (not the reference's): no virtual calls and no per-ray IOR evaluation. A lens is a flat table of
// LensIf records (144 B each) holding every per-interface constant the reference recomputes per ray
// (eta, 1/eta, R^2, 1/R ...; in the *_rgb variants all of them are ray-independent). The table lives in
// global memory and is staged into shared memory once per CTA; all lanes of a warp walk the same record,
// so the reads are broadcasts. All arithmetic is strict:: (see lrt_math.cuh) => miss masks are bit-exact.
#pragma once
#include "lrt_math.cuh"

namespace lrt {

enum : int { IF_SPHERE = 0, IF_PLANE = 1, IF_APERTURE = 2, IF_ASPHERE = 3 };
constexpr int kMaxInterfaces = 32;

struct alignas(16) LensIf {
    int kind; float z, elem_radius, elem_r2;
    float ior_left, ior_right, eta, rcp_eta;
    float curv_radius, center_z, curv_r2; int convex;
    float plane_param, c, K, r_scale;
    float ai[7]; float rcp_R;
    float gi[7]; float rcp_R2;
    float lim, k1, one_p_K, c2;
};
static_assert(sizeof(LensIf) == 144, "LensIf layout");

constexpr int kPupilSegments = 64;               // realisticlens.cpp:1772
constexpr int kPupilRaysLog2 = 20;               // 1024 * 1024 rays per segment (:1686)
struct alignas(16) LensSystem {
    int n; int kind;   // kind: 0 = realisticlens, 1 = perspective pinhole
    float rear_z, rear_radius;
    float film_z, near_clip, far_clip, shutter_open_time;
    float s2f_m00, s2f_m03, s2f_m11, s2f_m13;   // sample -> film (diagonal + translation)
    float to_world[12];                         // rows of the affine camera-to-world matrix
    // perspective (src/sensors/perspective.cpp): sample -> camera and image-plane normalisation
    float persp_s2c[16]; float persp_norm; int pupil_mode;   // pupil_mode: sample_exit_pupil (see pupil_bounds)
    const float *pupil_lut;   // -> pupil_bounds of the copy that lives in the same address space (host struct: host array, device
                              //    struct: device array); kernels that stage only the header + n interfaces read the LUT through it
    // film -> sample (diagonal + translation), film-plane normalisation 1/area and crop resolution: what
    // sample_direction needs (realisticlens.cpp:2190-2207,2352-2372)
    float f2s_m00, f2s_m03, f2s_m11, f2s_m13;
    float normalization, res_x, res_y, film_diagonal;
    LensIf ifs[kMaxInterfaces];
    // exit-pupil LUT (realisticlens.cpp:1769-1809): per radial film segment, (min.x, min.y, max.x, max.y) of the region of
    // the rear plane z = rear_z through which rays from that segment leave the lens; filled by lrt_sensor_set_exit_pupil
    float pupil_bounds[kPupilSegments][4];
};

struct LensRay { strict::V3 o, d; };

// ---- Dr.Jit sincos, single precision (drjit/math.h:76-176), strict
LRT_HD void sincos_drjit(float x, float &s_out, float &c_out) {
    using namespace strict;
    float xa = fabsf(x);
    int32_t j = (int32_t) mul(xa, 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    uint32_t sign_sin = ((uint32_t) j << 29) ^ f2u(x);
    uint32_t sign_cos = (uint32_t) (~(j - 2)) << 29;
    y = sub(sub(sub(xa, mul(y, 0.78515625f)), mul(y, 2.4187564849853515625e-4f)), mul(y, 3.77489497744594108e-8f));
    float z = sqr(y);
    if (xa == inf_f()) z = u2f(0xffffffffu);
    float z2 = sqr(z);
    float s = mul(fmadd(z2, -1.9515295891e-4f, fmadd(z, 8.3321608736e-3f, -1.6666654611e-1f)), z);
    float c = mul(fmadd(z2, 2.443315711809948e-5f, fmadd(z, -1.388731625493765e-3f, 4.166664568298827e-2f)), z);
    s = fmadd(s, y, y);
    c = fmadd(c, z, fmadd(z, -0.5f, 1.f));
    bool polymask = (j & 2) == 0;
    s_out = u2f(f2u(polymask ? s : c) ^ (sign_sin & 0x80000000u));
    c_out = u2f(f2u(polymask ? c : s) ^ (sign_cos & 0x80000000u));
}

// ---- warp.h:54-90
LRT_HD void square_to_disk_concentric(float u, float v, float &px, float &py) {
    using namespace strict;
    float x = fmsub(2.f, u, 1.f), y = fmsub(2.f, v, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f);
    bool q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = div(mul(mul(0.25f, kPi), rp), r);
    if (q13) phi = sub(mul(0.5f, kPi), phi);
    if (is_zero) phi = 0.f;
    float s, c; sincos_drjit(phi, s, c);
    px = mul(r, c); py = mul(r, s);
}

// ---- fresnel.h:35-84 (only what the lens needs: r for the TIR test, cos_theta_t, eta_ti)
LRT_HD void lens_fresnel(float cos_theta_i, float eta, float rcp_eta, float &r, float &cos_theta_t, float &eta_ti) {
    using namespace strict;
    bool outside = cos_theta_i >= 0.f;
    float eta_it = outside ? eta : rcp_eta;
    eta_ti = outside ? rcp_eta : eta;
    float ct2 = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), mul(eta_ti, eta_ti), 1.f);
    float ci = fabsf(cos_theta_i);
    float ct = sqrt(fmaxf(ct2, 0.f));
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float a_s = div(fnmadd(eta_it, ct, ci), fmadd(eta_it, ct, ci));
    float a_p = div(fnmadd(eta_it, ci, ct), fmadd(eta_it, ci, ct));
    r = mul(0.5f, add(sqr(a_s), sqr(a_p)));
    if (special) r = index_matched ? 0.f : 1.f;
    cos_theta_t = cos_theta_i >= 0.f ? -ct : ct;
}

struct LensHit { float t; strict::V3 p, n; };   // valid <=> t != inf

LRT_HD LensHit hit_none() { LensHit h; h.t = inf_f(); h.p = strict::v3(0, 0, 0); h.n = strict::v3(0, 0, 0); return h; }

LRT_HD LensHit hit_sphere(const LensIf &f, const LensRay &ray) {
    using namespace strict;
    LensHit si = hit_none();
    V3 center = v3(0.f, 0.f, f.center_z);
    V3 pcl = sub(center, ray.o);
    float center_proj = dot(pcl, ray.d);
    float perp = norm(sub(pcl, mul(ray.d, center_proj)));
    float disc = sub(f.curv_r2, sqr(perp));
    if (!(disc >= 0.f)) return si;
    float sq = sqrt(disc);
    float near_t = sub(center_proj, sq), far_t = add(center_proj, sq);
    if (!(far_t >= 0.f)) return si;
    float t = ((f.convex != 0) != (ray.d.z < 0.f)) ? (near_t >= 0.f ? near_t : inf_f()) : far_t;
    V3 p = at(ray.o, ray.d, t);
    V3 n = normalize(sub(p, center));
    if (!f.convex) n = neg(n);
    si.t = t; si.p = p; si.n = n;
    return si;
}

LRT_HD LensHit hit_plane(const LensIf &f, const LensRay &ray) {
    using namespace strict;
    LensHit si = hit_none();
    V3 n = v3(0.f, 0.f, -1.f);
    float n_dot_d = dot(n, ray.d);
    if (!(fabsf(n_dot_d) >= kEpsilon)) return si;
    float t = div(sub(f.plane_param, dot(n, ray.o)), n_dot_d);
    if (!(t >= 0.f)) return si;
    si.t = t; si.p = at(ray.o, ray.d, t); si.n = n;
    return si;
}

// sag / slope of the even asphere in unitless radius (realisticlens.cpp:680-728)
LRT_HD float asph_eval(const LensIf &f, float r2) {
    using namespace strict;
    float r2_ = mul(r2, f.rcp_R2);
    float z_ = 0.f;
#pragma unroll
    for (int i = 6; i >= 0; --i) z_ = fmadd(z_, r2_, f.ai[i]);
    z_ = mul(z_, sqr(r2_));
    float sqr_term = sub(1.f, mul(f.k1, r2_));
    float conic = mul(mul(f.c, r2_), rcp(add(1.f, sqrt(sqr_term))));
    z_ = add(z_, conic);
    return add(mul(-z_, f.elem_radius), f.z);
}
LRT_HD strict::V3 asph_normal(const LensIf &f, strict::V3 p) {
    using namespace strict;
    float rx = mul(p.x, f.rcp_R), ry = mul(p.y, f.rcp_R);
    float r2_ = add(mul(rx, rx), mul(ry, ry));
    if (r2_ >= 4.f * kEpsilon) { float s = rsqrt(r2_); rx = mul(rx, s); ry = mul(ry, s); } else { rx = 0.f; ry = 0.f; }
    float r_ = sqrt(r2_);
    float z_ = 0.f;
#pragma unroll
    for (int i = 6; i >= 0; --i) z_ = fmadd(z_, r2_, f.gi[i]);
    z_ = mul(z_, mul(r2_, r_));
    float cr = mul(f.c, r_);
    float sqr_term = sub(1.f, mul(f.one_p_K, sqr(cr)));
    z_ = add(z_, mul(cr, rsqrt(sqr_term)));
    float zg = -z_;
    return normalize(v3(mul(zg, rx), mul(zg, ry), -1.0f));
}

// core/math.h:360-402
LRT_HD bool solve_quadratic_strict(float a, float b, float c, float &x0, float &x1) {
    using namespace strict;
    bool linear_case = a == 0.f;
    bool valid_linear = linear_case && b != 0.f;
    x0 = x1 = div(-c, b);
    float discrim = fmsub(b, b, mul(mul(4.f, a), c));
    bool valid_quadratic = !linear_case && discrim >= 0.f;
    if (valid_quadratic) {
        float sd = sqrt(discrim);
        float temp = mul(-0.5f, add(b, copysignf(sd, b)));
        float x0p = div(temp, a), x1p = div(c, temp);
        x0 = fminf(x0p, x1p); x1 = fmaxf(x0p, x1p);
    }
    return valid_linear || valid_quadratic;
}

// realisticlens.cpp:770-818
LRT_HD bool asph_intersect_conic(const LensIf &f, const LensRay &ray, float &t_out) {
    using namespace strict;
    V3 o = sub(ray.o, v3(0.f, 0.f, f.z)), d = ray.d;
    float R = f.elem_radius;
    o = mul(o, f.rcp_R);
    float A = mul(f.c, add(1.f, mul(f.K, sqr(d.z))));
    float B = mul(2.f, sub(mul(f.c, add(dot(o, d), mul(mul(f.K, o.z), d.z))), d.z));
    float C = sub(mul(f.c, add(squared_norm(o), mul(f.K, sqr(o.z)))), mul(2.f, o.z));
    float t0, t1;
    bool valid = solve_quadratic_strict(A, B, C, t0, t1);
    if (!valid) { t_out = t0; return false; }
    t0 = mul(t0, R); t1 = mul(t1, R);
    float t0z = mul(f.c, sub(add(ray.o.z, mul(t0, ray.d.z)), f.z));
    float t1z = mul(f.c, sub(add(ray.o.z, mul(t1, ray.d.z)), f.z));
    bool t0v = t0z > 0.f && t0z <= f.lim, t1v = t1z > 0.f && t1z <= f.lim;
    if (!(t0v || t1v)) { t_out = t0; return false; }
    if (t0v && t1v) { t_out = t0 > 0.f ? t0 : t1; return true; }
    t_out = t0v ? t0 : t1; return true;
}

// realisticlens.cpp:574-645: Newton on tangent planes from the conic guess, at most 10 iterations
LRT_HD LensHit hit_asphere(const LensIf &f, const LensRay &ray) {
    using namespace strict;
    const float TOL = kEpsilon * 10.f;
    LensHit si = hit_none();
    float t;
    if (!asph_intersect_conic(f, ray, t)) t = mul(sub(f.z, ray.o.z), rcp(ray.d.z));
    V3 p = at(ray.o, ray.d, t);
    float r2 = add(sqr(p.x), sqr(p.y)), err = inf_f();
    int itr = 0; bool active = true;
    while (active) {
        float z_asph = asph_eval(f, r2);
        V3 plane_p = v3(p.x, p.y, z_asph);
        err = fabsf(sub(p.z, z_asph));
        V3 plane_n = asph_normal(f, p);
        t = div(dot(plane_n, sub(plane_p, ray.o)), dot(plane_n, ray.d));
        p = at(ray.o, ray.d, t);
        r2 = add(sqr(p.x), sqr(p.y));
        itr++;
        active = (err > TOL) && (itr < 10);
    }
    if (!(err < TOL)) return si;
    si.t = t; si.p = p; si.n = asph_normal(f, p);
    return si;
}

// LensInterface::compute_interaction (:269-297) / ApertureStop (:505-517).
// Returns the `active` mask of the reference; `ray` is advanced only when active.
template <bool kOwnEta> LRT_HD bool lens_interact_t(const LensIf &f, LensRay &ray, float eta_ray, float rcp_eta_ray) {
    using namespace strict;
    LensHit si = f.kind == IF_SPHERE ? hit_sphere(f, ray) : (f.kind == IF_ASPHERE ? hit_asphere(f, ray) : hit_plane(f, ray));
    bool active = si.t != inf_f();
    active = active && (add(sqr(si.p.x), sqr(si.p.y)) < f.elem_r2);
    if (f.kind == IF_APERTURE) {
        if (active) ray.o = si.p;
        return active;
    }
    V3 wi = neg(ray.d);
    float cos_theta_i = dot(wi, si.n);
    float r, ct, eti;
    if (kOwnEta) lens_fresnel(cos_theta_i, eta_ray, rcp_eta_ray, r, ct, eti);
    else lens_fresnel(cos_theta_i, f.eta, f.rcp_eta, r, ct, eti);
    active = active && (r <= kOneMinusEpsilon);
    // refract (fresnel.h:311-315): fmsub(m, fmadd(dot(wi,m), eta_ti, cos_theta_t), wi*eta_ti)
    float g = fmadd(cos_theta_i, eti, ct);
    V3 d_out = v3(sub(mul(si.n.x, g), mul(wi.x, eti)), sub(mul(si.n.y, g), mul(wi.y, eti)), sub(mul(si.n.z, g), mul(wi.z, eti)));
    if (active) { ray.o = si.p; ray.d = d_out; }
    return active;
}

LRT_HD bool lens_interact(const LensIf &f, LensRay &ray) { return lens_interact_t<false>(f, ray, 0.f, 0.f); }

// ---- the spectral variants' lens step (`*_spectral*`, the fork's dispersion mode): every ray carries a wavelength and
// DispersiveMaterial::compute_ior(ray) (:169-183) evaluates both media of every interface at 0.001 * wavelengths[0]
// (nm -> um): Cauchy A + B / w^2 or Sellmeier sqrt(1 + sum B_i w^2 / (w^2 - C_i)) (:226-240). mat = 7 floats
// [0 = Cauchy | 1 = Sellmeier, A, B, -, -, -, -] or [1, B1, B2, B3, C1, C2, C3]; mat_lr = per interface (left 7, right 7).
LRT_HD float lens_material_ior(const float *m, float w_um) {
    using namespace strict;
    const float w2 = sqr(w_um);
    if (m[0] == 0.f) return add(m[1], div(m[2], w2));
    float ior = 1.f;
    ior = add(ior, div(mul(m[1], w2), sub(w2, m[4])));
    ior = add(ior, div(mul(m[2], w2), sub(w2, m[5])));
    ior = add(ior, div(mul(m[3], w2), sub(w2, m[6])));
    return sqrt(ior);
}
LRT_HD bool lens_interact_wl(const LensIf &f, LensRay &ray, const float *mat_lr, float w_um) {
    // compute_interaction (:279-284): ext_ior = left, int_ior = right, eta = int / ext; fresnel() takes its own reciprocal
    const float ext_ior = lens_material_ior(mat_lr, w_um), int_ior = lens_material_ior(mat_lr + 7, w_um);
    const float eta = strict::div(int_ior, ext_ior);
    return lens_interact_t<true>(f, ray, eta, strict::rcp(eta));
}
LRT_HD bool lens_trace_wl(const LensIf *ifs, int n, LensRay &ray, const float *mat_lr, float wavelength_nm, bool from_world) {
    const float w_um = strict::mul(0.001f, wavelength_nm);
    if (from_world) { for (int i = n - 1; i >= 0; --i) if (!lens_interact_wl(ifs[i], ray, mat_lr + 14 * i, w_um)) return false; }
    else { for (int i = 0; i < n; ++i) if (!lens_interact_wl(ifs[i], ray, mat_lr + 14 * i, w_um)) return false; }
    return true;
}

LRT_HD bool lens_trace_from_film(const LensIf *ifs, int n, LensRay &ray) {
    for (int i = 0; i < n; ++i)
        if (!lens_interact(ifs[i], ray)) return false;
    return true;
}
LRT_HD bool lens_trace_from_world(const LensIf *ifs, int n, LensRay &ray) {
    for (int i = n - 1; i >= 0; --i)
        if (!lens_interact(ifs[i], ray)) return false;
    return true;
}

struct CameraRay { float3 o, d; float maxt; float weight; bool active; float3 d_local; };

// RealisticLensCamera::sample_ray (:2209-2292), rear-element sampling branch. (u,v): film sample in
// [0,1]^2 of the crop window, (a,b): aperture sample.
// RadicalInverse::eval (include/mitsuba/core/qmc.h:68-91), unscrambled, for the prime bases 2, 3, 5 (base_index 0..2):
// the digits of `index` mirrored at the radix point; the scale is a running product of the rounded reciprocal.
LRT_HD float radical_inverse(int base_index, uint64_t index) {
    const uint64_t divisor = base_index == 0 ? 2u : (base_index == 1 ? 3u : 5u);
    const float recip = strict::div(1.f, (float) divisor);
    uint64_t value = 0; float factor = 1.f;
    while (index != 0) {
        const uint64_t next = index / divisor;
        factor = strict::mul(factor, recip);
        value = (value - next) * divisor + index;
        index = next;
    }
    return fminf(u2f(0x3f7fffffu), strict::mul((float) value, factor));
}

// dr::lerp (drjit/array_router.h:332): fmadd(b, t, fnmadd(a, t, a)), every step rounded
LRT_HD float lerp_drjit(float a, float b, float t) { return strict::add(strict::mul(b, t), strict::sub(a, strict::mul(a, t))); }

// sample_exit_pupil (realisticlens.cpp:1812-1845): a point of the rear plane inside the LUT box of the film point's radial
// segment, rotated to the film point's azimuth; `bounds_area` = area of that box
LRT_HD strict::V3 lens_sample_exit_pupil(const LensSystem &L, strict::V3 film_p, float a, float b, float &bounds_area) {
    using namespace strict;
    const float r_film = sqrt(add(sqr(film_p.x), sqr(film_p.y)));
    int idx = (int) floorf(mul(div(r_film, L.film_diagonal), (float) kPupilSegments));
    idx = idx < 0 ? 0 : (idx > kPupilSegments - 1 ? kPupilSegments - 1 : idx);
    const float *lut = L.pupil_lut + 4 * idx;
    const float minx = lut[0], miny = lut[1], maxx = lut[2], maxy = lut[3];
    bounds_area = mul(sub(maxx, minx), sub(maxy, miny));
    const float px = lerp_drjit(minx, maxx, a), py = lerp_drjit(miny, maxy, b);
    const float sin_t = r_film > 0.f ? mul(film_p.y, rcp(r_film)) : 0.f, cos_t = r_film > 0.f ? mul(film_p.x, rcp(r_film)) : 1.f;
    return v3(sub(mul(cos_t, px), mul(sin_t, py)), add(mul(sin_t, px), mul(cos_t, py)), L.rear_z);
}

struct TraceFromFilm { const LensIf *ifs; int n; LRT_HD bool operator()(LensRay &ray) const { return lens_trace_from_film(ifs, n, ray); } };
struct TraceFromFilmWl {
    const LensIf *ifs; int n; const float *mat_lr; float wavelength_nm;
    LRT_HD bool operator()(LensRay &ray) const { return lens_trace_wl(ifs, n, ray, mat_lr, wavelength_nm, false); }
};

template <class Trace> LRT_HD CameraRay lens_sample_ray_with(const LensSystem &L, float u, float v, float a, float b, const Trace &trace) {
    using namespace strict;
    V3 film_p = v3(add(mul(L.s2f_m00, u), L.s2f_m03), add(mul(L.s2f_m11, v), L.s2f_m13), 0.f);
    V3 ap; float bounds_area = 0.f;
    if (L.pupil_mode) ap = lens_sample_exit_pupil(L, film_p, a, b, bounds_area);
    else {
        float dx, dy; square_to_disk_concentric(a, b, dx, dy);
        ap = v3(mul(L.rear_radius, dx), mul(L.rear_radius, dy), L.rear_z);
    }
    LensRay ray; ray.o = film_p; ray.d = normalize(sub(ap, film_p));
    bool active = trace(ray);
    V3 dl = ray.d, o = ray.o;
    o.z = add(o.z, L.film_z);
    const float *m = L.to_world;
    V3 ow = v3(fmadd(m[2], o.z, fmadd(m[1], o.y, fmadd(m[0], o.x, m[3]))),
               fmadd(m[6], o.z, fmadd(m[5], o.y, fmadd(m[4], o.x, m[7]))),
               fmadd(m[10], o.z, fmadd(m[9], o.y, fmadd(m[8], o.x, m[11]))));
    V3 dw = v3(fmadd(m[2], dl.z, fmadd(m[1], dl.y, mul(m[0], dl.x))),
               fmadd(m[6], dl.z, fmadd(m[5], dl.y, mul(m[4], dl.x))),
               fmadd(m[10], dl.z, fmadd(m[9], dl.y, mul(m[8], dl.x))));
    float inv_z = rcp(dl.z);
    float near_t = mul(L.near_clip, inv_z), far_t = mul(L.far_clip, inv_z);
    ow = at(ow, dw, near_t);
    CameraRay r;
    r.o = f3(ow); r.d = f3(dw); r.maxt = sub(far_t, near_t);
    // weight (:2279-2289): cos^4 of the exit direction, or in pupil mode shutter_open_time * box area / rear_z^2
    // (the reference's shutter_open_time() is 0 unless shutter_open/close are set: its pupil-mode images are then black)
    const float w = L.pupil_mode ? mul(mul(L.shutter_open_time, bounds_area), rcp(sqr(L.rear_z))) : sqr(sqr(dl.z));
    r.weight = active ? w : 0.f;
    r.active = active; r.d_local = f3(dl);
    return r;
}
LRT_HD CameraRay lens_sample_ray(const LensSystem &L, const LensIf *ifs, float u, float v, float a, float b) {
    return lens_sample_ray_with(L, u, v, a, b, TraceFromFilm { ifs, L.n });
}
// the spectral variants' sample_ray: the geometric part (the caller multiplies the weight by sample_wavelengths()' own)
LRT_HD CameraRay lens_sample_ray_wl(const LensSystem &L, const LensIf *ifs, const float *mat_lr, float wavelength_nm, float u, float v, float a, float b) {
    return lens_sample_ray_with(L, u, v, a, b, TraceFromFilmWl { ifs, L.n, mat_lr, wavelength_nm });
}


// RealisticLensCamera::sample_direction (:2301-2394) in CAMERA space (the caller applies to_world^-1 to the
// reference point and to_world to p / d): sample the front-element disk, trace front -> rear, project the ray
// to the film plane z = 0. Returns the `valid` mask; uv in pixels of the crop window (ds.uv).
struct DirectionSample { float uv_x, uv_y; strict::V3 p, d; float dist, pdf, weight; bool valid; LensRay ray_out; };

// `origin_offset` = 0 reproduces the reference (the ray starts at ref_p). A positive value starts the SAME ray
// origin_offset metres in front of the sampled front-plane point instead: from a distant reference point the
// float32 products t*d cancel against o (0.3 px of noise from 1 km) and the asphere's absolute Newton tolerance
// (10 eps metres) is unreachable, so the lens-design path uses a nearby origin.
LRT_HD DirectionSample lens_sample_direction(const LensSystem &L, const LensIf *ifs, strict::V3 ref_p, float a, float b,
                                             float origin_offset = 0.f) {
    using namespace strict;
    DirectionSample ds; ds.uv_x = ds.uv_y = 0.f; ds.p = v3(0, 0, 0); ds.d = v3(0, 0, 0); ds.dist = 0.f; ds.pdf = 0.f; ds.weight = 0.f;
    ds.valid = false; ds.ray_out.o = v3(0, 0, 0); ds.ray_out.d = v3(0, 0, 0);
    bool active = ref_p.z >= L.near_clip && ref_p.z <= L.far_clip;
    if (!active) return ds;
    const LensIf &front = ifs[L.n - 1];
    float front_radius = mul(1.0f, front.elem_radius), front_z = front.z;
    float dx, dy; square_to_disk_concentric(a, b, dx, dy);
    V3 aperture_p = v3(mul(dx, front_radius), mul(dy, front_radius), front_z);
    V3 dir = sub(ref_p, aperture_p);
    float dist = norm(dir), inv_dist = rcp(dist);
    dir = mul(dir, inv_dist);
    LensRay ray; ray.o = origin_offset > 0.f ? at(aperture_p, dir, origin_offset) : ref_p; ray.d = neg(dir);
    bool valid = lens_trace_from_world(ifs, L.n, ray);
    ds.ray_out = ray;
    if (!valid) return ds;
    float t = div(-ray.o.z, ray.d.z);
    V3 pf = at(ray.o, ray.d, t);
    // Transform4f::transform_affine column chain (transform.h:104-111); film_to_sample is diagonal + translation
    float sx = fmadd(0.f, pf.z, fmadd(0.f, pf.y, fmadd(L.f2s_m00, pf.x, L.f2s_m03)));
    float sy = fmadd(0.f, pf.z, fmadd(L.f2s_m11, pf.y, fmadd(0.f, pf.x, L.f2s_m13)));
    valid = valid && sx >= 0.f && sy >= 0.f && sx <= 1.f && sy <= 1.f;
    float ct_film = -ray.d.z, inv_ct = rcp(ct_film);
    float value = valid ? mul(mul(L.normalization, sqr(sqr(inv_ct))), sqr(L.rear_z)) : 0.f;
    if (!valid) return ds;
    ds.valid = true;
    ds.uv_x = mul(sx, L.res_x); ds.uv_y = mul(sy, L.res_y);
    ds.p = aperture_p;
    ds.d = mul(sub(aperture_p, ref_p), inv_dist);
    ds.dist = dist;
    float aperture_pdf = rcp(mul(kPi, sqr(front_radius)));
    float ct_ref = dir.z;
    ds.pdf = mul(mul(mul(aperture_pdf, dist), dist), rcp(ct_ref));
    ds.weight = mul(mul(mul(value, inv_dist), inv_dist), ct_ref);
    return ds;
}

// perspective pinhole (src/sensors/perspective.cpp sample_ray): near_p = s2c * (u,v,0), d = normalize(near_p)
LRT_HD CameraRay persp_sample_ray(const LensSystem &L, float u, float v) {
    const float *m = L.persp_s2c;
    float x = m[0] * u + m[1] * v + m[3], y = m[4] * u + m[5] * v + m[7], z = m[8] * u + m[9] * v + m[11],
          w = m[12] * u + m[13] * v + m[15];
    float iw = 1.f / w;
    float3 d = normalize(make_float3(x * iw, y * iw, z * iw));
    const float *t = L.to_world;
    float3 o = make_float3(t[3], t[7], t[11]);
    float3 dw = make_float3(t[0] * d.x + t[1] * d.y + t[2] * d.z, t[4] * d.x + t[5] * d.y + t[6] * d.z,
                            t[8] * d.x + t[9] * d.y + t[10] * d.z);
    float inv_z = 1.f / d.z;
    float near_t = L.near_clip * inv_z, far_t = L.far_clip * inv_z;
    CameraRay r;
    r.o = o + dw * near_t; r.d = dw; r.maxt = far_t - near_t; r.weight = 1.f; r.active = true; r.d_local = d;
    return r;
}

LRT_HD CameraRay sensor_sample_ray(const LensSystem &L, const LensIf *ifs, float u, float v, float a, float b) {
    return L.kind == 0 ? lens_sample_ray(L, ifs, u, v, a, b) : persp_sample_ray(L, u, v);
}

}  // namespace lrt
