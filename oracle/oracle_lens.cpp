// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). Lens step restatement:
// src/sensors/realisticlens.cpp (classes :147-819, camera :824-2466) and the small maths it calls.
#include "oracle.h"
#include <cstring>
#include <cstdio>

namespace orc {

static inline float bits2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline uint32_t f2bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

// ext/drjit/include/drjit/math.h:76-176 — CEPHES-style joint sin/cos, single precision.
// estrin(z, c0, c1, c2) = fmadd(z*z, c2, fmadd(z, c1, c0))  (math.h:27-41)
void sincos(float x, float &s_out, float &c_out) {
    float xa = std::fabs(x);
    int32_t j = (int32_t) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    uint32_t sign_sin = ((uint32_t) j << 29) ^ f2bits(x);
    uint32_t sign_cos = (uint32_t) (~(j - 2)) << 29;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == Infinity) z = bits2f(0xffffffffu);
    float s = fmadd(z * z, -1.9515295891e-4f, fmadd(z, 8.3321608736e-3f, -1.6666654611e-1f)) * z;
    float c = fmadd(z * z, 2.443315711809948e-5f,
                    fmadd(z, -1.388731625493765e-3f, 4.166664568298827e-2f)) * z;
    s = fmadd(s, y, y);
    c = fmadd(c, z, fmadd(z, -0.5f, 1.f));
    bool polymask = (j & 2) == 0;
    // mulsign(v, sign) = v xor (sign & 0x80000000)  (array_router.h:380-387)
    s_out = bits2f(f2bits(polymask ? s : c) ^ (sign_sin & 0x80000000u));
    c_out = bits2f(f2bits(polymask ? c : s) ^ (sign_cos & 0x80000000u));
}

// ext/drjit/include/drjit/math.h:519-594 — atan2, single precision: minimax polynomial in (min/max)^2, Estrin's scheme
// (math.h:26-41: pairs combined with fmadd, then the same on x^2, x^4)
float atan2_drjit(float y, float x) {
    float abs_x = std::fabs(x), abs_y = std::fabs(y), min_val = std::fmin(abs_y, abs_x), max_val = std::fmax(abs_x, abs_y);
    float scaled_min = min_val / max_val, z = sqr(scaled_min);
    const float c0 = 0.99999934166683966009f, c1 = -0.33326497518773606976f, c2 = 0.19881342388439013552f, c3 = -0.13486708938456973185f,
                c4 = 0.083863120428809689910f, c5 = -0.037006525670417265220f, c6 = 0.0078613793713198150252f;
    float a0 = fmadd(z, c1, c0), a1 = fmadd(z, c3, c2), a2 = fmadd(z, c5, c4), a3 = c6, z2 = sqr(z);
    float b0 = fmadd(z2, a1, a0), b1 = fmadd(z2, a3, a2);
    float t = fmadd(sqr(z2), b1, b0);
    t = t * scaled_min;
    if (abs_y > abs_x) t = Pi * .5f - t;
    if (x < 0.f) t = Pi - t;
    float r = y < 0.f ? -t : t;
    return max_val != 0.f ? r : 0.f;
}
// math.h:366-412 — asin, single precision (CEPHES)
float asin_drjit(float x) {
    float xa = std::fabs(x), x2 = sqr(x);
    bool big = xa > .5f;
    float x1 = .5f * (1.f - xa), x3 = big ? x1 : x2, x4 = big ? std::sqrt(x1) : xa;
    const float c0 = 1.6666752422e-1f, c1 = 7.4953002686e-2f, c2 = 4.5470025998e-2f, c3 = 2.4181311049e-2f, c4 = 4.2163199048e-2f;
    float a0 = fmadd(x3, c1, c0), a1 = fmadd(x3, c3, c2), a2 = c4, y2 = sqr(x3);
    float b0 = fmadd(y2, a1, a0), b1 = a2;
    float z1 = fmadd(sqr(y2), b1, b0);
    z1 = fmadd(z1, x3 * x4, x4);
    float r = big ? (.5f * Pi) - (z1 + z1) : z1;
    return std::copysign(r, x);
}

// include/mitsuba/core/warp.h:54-90
V2 square_to_uniform_disk_concentric(V2 u) {
    float x = fmsub(2.f, u.x, 1.f), y = fmsub(2.f, u.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f);
    bool q13 = std::fabs(x) < std::fabs(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * Pi * rp / r;
    if (q13) phi = 0.5f * Pi - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincos(phi, s, c);
    return { r * c, r * s };
}

// include/mitsuba/core/warp.h:241-252 (square_to_cosine_hemisphere)
V3 square_to_cosine_hemisphere(V2 u) {
    V2 p = square_to_uniform_disk_concentric(u);
    float z = std::sqrt(std::fmax(1.f - (p.x * p.x + p.y * p.y), 0.f));   // safe_sqrt(1 - |p|^2)
    return { p.x, p.y, z };
}

// include/mitsuba/render/fresnel.h:35-84
void fresnel(float cos_theta_i, float eta, float &r, float &cos_theta_t, float &eta_it, float &eta_ti) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = rcp(eta);
    eta_it = outside ? eta : rcp_eta;
    eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fnmadd(fnmadd(cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = std::fabs(cos_theta_i);
    float ct = std::sqrt(std::fmax(cos_theta_t_sqr, 0.f));
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fnmadd(eta_it, ct, ci) / fmadd(eta_it, ct, ci);
    float a_p = fnmadd(eta_it, ci, ct) / fmadd(eta_it, ci, ct);
    r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = r_sc;
    cos_theta_t = cos_theta_i >= 0.f ? -ct : ct;   // mulsign_neg (array_router.h:389-396)
}

// include/mitsuba/render/fresnel.h:311-315: fmsub(m, fmadd(dot(wi,m), eta_ti, cos_theta_t), wi*eta_ti)
V3 refract(V3 wi, V3 m, float cos_theta_t, float eta_ti) {
    float f = fmadd(dot(wi, m), eta_ti, cos_theta_t);
    return { m.x * f - wi.x * eta_ti, m.y * f - wi.y * eta_ti, m.z * f - wi.z * eta_ti };
}

// include/mitsuba/core/math.h:360-402
bool solve_quadratic(float a, float b, float c, float &x0, float &x1) {
    bool linear_case = a == 0.f;
    bool valid_linear = linear_case && b != 0.f;
    x0 = x1 = -c / b;
    float discrim = fmsub(b, b, 4.f * a * c);
    bool valid_quadratic = !linear_case && discrim >= 0.f;
    if (valid_quadratic) {
        float sqrt_discrim = std::sqrt(discrim);
        float temp = -0.5f * (b + std::copysign(sqrt_discrim, b));
        float x0p = temp / a, x1p = c / temp;
        float x0m = std::fmin(x0p, x1p), x1m = std::fmax(x0p, x1p);
        x0 = linear_case ? x0 : x0m;
        x1 = linear_case ? x0 : x1m;
    }
    return valid_linear || valid_quadratic;
}


// ================================================================================ lens classes
struct Hit { float t; V3 p, n; bool valid; };   // Interaction3f: valid <=> t != inf (zeros<> sets t=inf)

void Interface::offset_along_axis(float delta) {          // realisticlens.cpp:265-267,404-407,475-478
    z += delta;
    if (kind == SPHERE) center_z = z + curv_radius;
    if (kind == PLANE || kind == APERTURE) plane_param = -1.0f * z;   // m_normal.z() * z
}

// realisticlens.cpp:359-402 SpheroidLens::intersect/normal
static Hit intersect_sphere(const Interface &f, const Ray &ray) {
    Hit si { Infinity, { 0, 0, 0 }, { 0, 0, 0 }, false };
    V3 center { 0.f, 0.f, f.center_z };
    V3 pcl = center - ray.o;
    float center_proj = dot(pcl, ray.d);
    float perp = norm(pcl - center_proj * ray.d);
    float disc = sqr(f.curv_radius) - sqr(perp);
    if (!(disc >= 0.f)) return si;
    float sq = std::sqrt(disc);
    float near_t = center_proj - sq, far_t = center_proj + sq;
    if (!(far_t >= 0.f)) return si;
    float t = (f.convex != (ray.d.z < 0.f)) ? (near_t >= 0.f ? near_t : Infinity) : far_t;
    V3 p = ray.o + ray.d * t;           // ray(t) = fmadd(d, t, o)  (ray.h operator())
    V3 n = normalize(p - center);
    if (!f.convex) n = -n;
    si.t = t; si.p = p; si.n = n; si.valid = t != Infinity;
    return si;
}

// realisticlens.cpp:444-473 PlanoLens::intersect (normal = (0,0,-1))
static Hit intersect_plane(const Interface &f, const Ray &ray) {
    Hit si { Infinity, { 0, 0, 0 }, { 0, 0, 0 }, false };
    V3 n { 0.f, 0.f, -1.f };
    float n_dot_d = dot(n, ray.d);
    if (!(std::fabs(n_dot_d) >= Epsilon)) return si;
    float t = (f.plane_param - dot(n, ray.o)) / n_dot_d;
    if (!(t >= 0.f)) return si;
    si.t = t; si.p = ray.o + ray.d * t; si.n = n; si.valid = t != Infinity;
    return si;
}

// realisticlens.cpp:680-692
static float asph_conic(const Interface &f, float r2_) {
    float sqr_term = 1.f - (1.f + f.K) * sqr(f.c) * r2_;
    return f.c * r2_ * rcp(1.f + std::sqrt(sqr_term));
}
static float asph_conic_grad(const Interface &f, float r_) {
    float cr = f.c * r_;
    float sqr_term = 1.f - (1.f + f.K) * sqr(cr);
    return cr * rsqrt(sqr_term);
}
// realisticlens.cpp:695-704
static float eval_asph(const Interface &f, float r2) {
    float r2_ = r2 * rcp(sqr(f.elem_radius));
    float z_ = 0.f;
    for (int i = 6; i >= 0; --i) z_ = fmadd(z_, r2_, f.ai[i]);
    z_ *= sqr(r2_);
    z_ += asph_conic(f, r2_);
    return -z_ * f.elem_radius + f.z;
}
// realisticlens.cpp:719-728
static float asph_grad_unitless(const Interface &f, float r2_) {
    float r_ = std::sqrt(r2_);
    float z_ = 0.f;
    for (int i = 6; i >= 0; --i) z_ = fmadd(z_, r2_, (2.f * i + 4.f) * f.ai[i]);
    z_ *= r2_ * r_;
    z_ += asph_conic_grad(f, r_);
    return -z_;
}
// realisticlens.cpp:647-656
static V3 asph_normal(const Interface &f, V3 p) {
    float rx = p.x * rcp(f.elem_radius), ry = p.y * rcp(f.elem_radius);
    float r2_ = rx * rx + ry * ry;
    if (r2_ >= 4.f * Epsilon) { float s = rsqrt(r2_); rx *= s; ry *= s; } else { rx = 0.f; ry = 0.f; }
    float zg = asph_grad_unitless(f, r2_);
    return normalize(V3 { zg * rx, zg * ry, -1.0f });
}
// realisticlens.cpp:770-818
static bool intersect_conic(const Interface &f, const Ray &ray, float &t_out) {
    V3 o = ray.o - V3 { 0.f, 0.f, f.z }, d = ray.d;
    float R = f.elem_radius;
    o = o * rcp(R);
    float A = f.c * (1.f + f.K * sqr(d.z)),
          B = 2.f * (f.c * (dot(o, d) + f.K * o.z * d.z) - d.z),
          C = f.c * (squared_norm(o) + f.K * sqr(o.z)) - 2.f * o.z;
    float t0, t1;
    bool valid = solve_quadratic(A, B, C, t0, t1);
    if (!valid) { t_out = t0; return false; }
    t0 *= R; t1 *= R;
    float t0z = f.c * (ray.o.z + t0 * ray.d.z - f.z), t1z = f.c * (ray.o.z + t1 * ray.d.z - f.z);
    bool t0v = valid && t0z > 0.f, t1v = valid && t1z > 0.f;
    float lim = f.K > -1.0f ? R * rcp(1.0f + f.K) : Infinity;
    t0v = t0v && t0z <= lim; t1v = t1v && t1z <= lim;
    valid = valid && (t0v || t1v);
    if (!valid) { t_out = t0; return false; }
    if (t0v && t1v) { t_out = t0 > 0.f ? t0 : t1; return true; }
    t_out = t0v ? t0 : t1; return true;
}
// realisticlens.cpp:574-645 AsphericalLens::intersect (Newton on tangent planes, <= 10 its)
static Hit intersect_asphere(const Interface &f, const Ray &ray) {
    const float TOL = Epsilon * 10.f;
    Hit si { Infinity, { 0, 0, 0 }, { 0, 0, 0 }, false };
    float t; bool valid = intersect_conic(f, ray, t);
    if (!valid) t = (f.z - ray.o.z) * rcp(ray.d.z);
    V3 p = ray.o + ray.d * t;
    float r2 = sqr(p.x) + sqr(p.y), err = Infinity;
    uint32_t itr = 0; bool active = true;
    while (active) {
        float z_asph = eval_asph(f, r2);
        V3 plane_p { p.x, p.y, z_asph };
        err = std::fabs(p.z - plane_p.z);
        V3 plane_n = asph_normal(f, p);
        t = dot(plane_n, plane_p - ray.o) / dot(plane_n, ray.d);
        p = ray.o + ray.d * t;
        r2 = sqr(p.x) + sqr(p.y);
        itr++;
        active = active && (err > TOL) && (itr < 10);
    }
    if (!(err < TOL)) return si;
    // `active &= t > 0` at :637 is dead code in the reference: si.t is set regardless
    si.t = t; si.p = ray.o + ray.d * t; si.n = asph_normal(f, si.p); si.valid = t != Infinity;
    return si;
}

// realisticlens.cpp:269-297 LensInterface::compute_interaction / :505-517 ApertureStop
bool interact(const Interface &f, const Ray &ray, Ray &next) {
    Hit si = f.kind == SPHERE ? intersect_sphere(f, ray)
           : f.kind == ASPHERE ? intersect_asphere(f, ray) : intersect_plane(f, ray);
    bool active = si.valid;
    active = active && (sqr(si.p.x) + sqr(si.p.y)) < sqr(f.elem_radius);
    if (f.kind == APERTURE) {
        if (active) { next.o = si.p; next.d = ray.d; next.maxt = Largest; }
        return active;
    }
    float cos_theta_i = dot(-ray.d, si.n);
    float eta = f.ior_right / f.ior_left;
    float r, ct, eit, eti;
    fresnel(cos_theta_i, eta, r, ct, eit, eti);
    active = active && (r <= OneMinusEpsilon);
    V3 d_out = refract(-ray.d, si.n, ct, eti);
    if (active) { next.o = si.p; next.d = d_out; next.maxt = Largest; }
    return active;
}

// realisticlens.cpp:1654-1683
bool Camera::trace_from_film(Ray ray, Ray &out) const {
    bool active = true; Ray curr = ray;
    for (size_t i = 0; i < ifs.size(); ++i) {
        Ray nxt = curr;
        active = interact(ifs[i], curr, nxt) && active;
        if (!active) break;
        curr = nxt;
    }
    out = curr; return active;
}
// realisticlens.cpp:1960-1984
bool Camera::trace_from_world(Ray ray, Ray &out) const {
    bool active = true; Ray curr = ray;
    for (int i = (int) ifs.size() - 1; i >= 0; --i) {
        Ray nxt = curr;
        active = interact(ifs[i], curr, nxt) && active;
        if (!active) break;
        curr = nxt;
    }
    out = curr; return active;
}

}  // namespace orc

// ================================================================================ camera
namespace orc {

// dr::pow(x, int) -> detail::powi, square-and-multiply (drjit/math.h:977-992)
static float powi(float x, int y) {
    uint32_t n = (uint32_t) (y < 0 ? -y : y); float result = 1.f;
    while (n) { if (n & 1) result *= x; x *= x; n >>= 1; }
    return y >= 0 ? result : rcp(result);
}

// realisticlens.cpp:226-240 at the nominal wavelength 0.5893 um (:169-173, *_rgb variants)
// mat7 = [model (0 Cauchy, 1 Sellmeier), p0..p5]
float material_ior(const float *m, float wavelength) {
    if (m[0] == 0.f) return m[1] + m[2] / sqr(wavelength);
    float w2 = sqr(wavelength), ior = 1.f;
    ior += m[1] * w2 / (w2 - m[4]);
    ior += m[2] * w2 / (w2 - m[5]);
    ior += m[3] * w2 / (w2 - m[6]);
    return std::sqrt(ior);
}

Interface make_sphere(float curv_radius, float elem_radius, float z, float il, float ir) {
    Interface f; f.kind = SPHERE; f.z = z; f.elem_radius = elem_radius; f.curv_radius = curv_radius;
    f.center_z = z + curv_radius; f.convex = curv_radius > 0.f; f.ior_left = il; f.ior_right = ir;
    return f;                                             // :349-356
}
Interface make_plane(int kind, float elem_radius, float z, float il, float ir) {
    Interface f; f.kind = kind; f.z = z; f.elem_radius = elem_radius; f.plane_param = -1.0f * z;
    f.ior_left = il; f.ior_right = ir; return f;          // :437-442, :499-501
}
Interface make_asphere(float curv_radius_mm, float kappa, float elem_radius, float z, const float *ai,
                       float il, float ir) {              // :557-572
    Interface f; f.kind = ASPHERE; f.z = z; f.elem_radius = elem_radius; f.K = kappa;
    f.r_scale = 1.0f / 0.001f * elem_radius;
    f.c = f.r_scale * rcp(curv_radius_mm);
    for (int i = 0; i < 7; ++i) f.ai[i] = powi(f.r_scale, 2 * i + 3) * ai[i];
    f.ior_left = il; f.ior_right = ir; return f;
}

// realisticlens.cpp:1568-1651 build_lens_from_data: tables front->rear in mm, mats has n+1 rows
void build_from_prescription(Camera &cam, int n, const int *kinds, const float *thick, const float *curv,
                             const float *elemr, const float *kappa, const float *ai7, const float *mats7,
                             float scale, float stop_ratio) {
    float z_pos = 0.f;
    for (int i = n - 1; i >= 0; --i) {
        float elem_radius = elemr[i];
        z_pos += thick[i];
        float il = material_ior(mats7 + 7 * (i + 1), 0.5893f), ir = material_ior(mats7 + 7 * i, 0.5893f);
        switch (kinds[i]) {
            case APERTURE:
                cam.ifs.push_back(make_plane(APERTURE, 0.001f * scale * elem_radius * stop_ratio,
                                             0.001f * scale * z_pos, il, il));
                break;
            case PLANE:
                cam.ifs.push_back(make_plane(PLANE, 0.001f * scale * elem_radius, 0.001f * scale * z_pos, il, ir));
                break;
            case ASPHERE: {
                float a[7];
                for (int j = 0; j < 7; ++j) a[j] = ai7[7 * i + j] / powi(scale, 2 * j + 3);
                cam.ifs.push_back(make_asphere(scale * curv[i], kappa[i], 0.001f * scale * elem_radius,
                                               0.001f * scale * z_pos, a, il, ir));
                break;
            }
            default:
                cam.ifs.push_back(make_sphere(-0.001f * scale * curv[i], 0.001f * scale * elem_radius,
                                              0.001f * scale * z_pos, il, ir));
        }
    }
}

// realisticlens.cpp:2107-2113
static void cardinal_points(const Ray &start, const Ray &end, float &z_p, float &z_f) {
    float t_focus = -end.o.x / end.d.x;
    z_f = fmadd(end.d.z, t_focus, end.o.z);
    float t_plane = (start.o.x - end.o.x) / end.d.x;
    z_p = fmadd(end.d.z, t_plane, end.o.z);
}

// realisticlens.cpp:2115-2171; returns false when a paraxial ray is blocked (reference throws)
bool focus_thick_lens(const Camera &cam, float focus_distance, float &delta) {
    float r = 0.001f * cam.film_half_x;
    Ray obj { { r, 0.f, cam.ifs.back().z + 1.f }, { 0.f, 0.f, -1.f } }, obj_end;
    if (!cam.trace_from_world(obj, obj_end)) return false;
    float obj_plane, obj_focus; cardinal_points(obj, obj_end, obj_plane, obj_focus);
    float p_img = obj_plane, f_img = obj_plane - obj_focus;
    Ray img { { r, 0.f, cam.ifs.front().z - 1.f }, { 0.f, 0.f, 1.f } }, img_end;
    if (!cam.trace_from_film(img, img_end)) return false;
    float img_plane, img_focus; cardinal_points(img, img_end, img_plane, img_focus);
    float p_obj = img_plane;
    float tmp = std::fmax(focus_distance, 4.01f * f_img) - p_obj;
    delta = 0.5f * (p_img - tmp + std::sqrt(sqr(p_img + tmp) - 4 * f_img * (p_img + tmp)));
    return true;
}

// film.h:193-198 get_physical_size + sensor.h:305-342 film_to_crop_transform + :2190-2207
void setup_film(Camera &cam, int w, int h, int cw, int ch, int cx, int cy, float diagonal_m) {
    float aspect = (float) h / (float) w;
    float x = std::sqrt(diagonal_m * diagonal_m / (1.f + aspect * aspect));
    float y = aspect * x;
    cam.film_half_x = x / 2; cam.film_half_y = y / 2;
    float px = cam.film_half_x, py = cam.film_half_y;
    cam.film_diagonal = std::sqrt(px * px + py * py);   // realisticlens.cpp:838 m_film_diagonal = norm(get_physical_size())
    float relx = (float) cw / (float) w, rely = (float) ch / (float) h;
    float offx = (float) cx / (float) w, offy = (float) cy / (float) h;
    // forward: S(1/rel) * T(-off) * S(0.5/phys) * T(phys); inverse composed from the parts' inverses
    float ax = rcp(1.f / relx), ay = rcp(1.f / rely);
    float cxs = rcp(0.5f / px), cys = rcp(0.5f / py);
    float *m = cam.sample_to_film;
    for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.f : 0.f;
    m[0] = ax * cxs; m[5] = ay * cys;
    m[3] = offx * cxs + (-px); m[7] = offy * cys + (-py);
    float *g = cam.film_to_sample;
    for (int i = 0; i < 16; ++i) g[i] = (i % 5 == 0) ? 1.f : 0.f;
    float fx = (1.f / relx) * (0.5f / px), fy = (1.f / rely) * (0.5f / py);
    g[0] = fx; g[5] = fy;
    g[3] = fx * px + (1.f / relx) * (-offx); g[7] = fy * py + (1.f / rely) * (-offy);
    // image rect / normalization (:2197-2203)
    float x0 = m[3], y0 = m[7], x1 = m[0] + m[3], y1 = m[5] + m[7];
    float minx = std::fmin(x0, x1), maxx = std::fmax(x0, x1), miny = std::fmin(y0, y1), maxy = std::fmax(y0, y1);
    cam.normalization = 1.f / ((maxx - minx) * (maxy - miny));
    cam.res_x = (float) cw; cam.res_y = (float) ch;
}

// realisticlens.cpp:898-912
void finish_camera(Camera &cam, bool fine_focus, float object_distance) {
    if (fine_focus) {
        float delta;
        if (!focus_thick_lens(cam, object_distance, delta)) delta = std::numeric_limits<float>::quiet_NaN();
        cam.focus_delta = delta;
        for (auto &f : cam.ifs) f.offset_along_axis(std::isnan(delta) ? 0.0f : -delta);
    }
    cam.rear_z = cam.ifs.front().z; cam.rear_radius = cam.ifs.front().elem_radius;
}

static V3 xform_point(const float *m, V3 p) {     // transform.h:104-111 (column fmadd chain)
    V3 r { m[3], m[7], m[11] };
    r = { fmadd(m[0], p.x, r.x), fmadd(m[4], p.x, r.y), fmadd(m[8], p.x, r.z) };
    r = { fmadd(m[1], p.y, r.x), fmadd(m[5], p.y, r.y), fmadd(m[9], p.y, r.z) };
    r = { fmadd(m[2], p.z, r.x), fmadd(m[6], p.z, r.y), fmadd(m[10], p.z, r.z) };
    return r;
}
static V3 xform_vector(const float *m, V3 v) {    // transform.h:132-140
    V3 r { m[0] * v.x, m[4] * v.x, m[8] * v.x };
    r = { fmadd(m[1], v.y, r.x), fmadd(m[5], v.y, r.y), fmadd(m[9], v.y, r.z) };
    r = { fmadd(m[2], v.z, r.x), fmadd(m[6], v.z, r.y), fmadd(m[10], v.z, r.z) };
    return r;
}

// RadicalInverse::eval (include/mitsuba/core/qmc.h:68-91), unscrambled; base table as src/core/qmc.cpp:46-52 fills it
// (recip = 1.f / (float) prime). base_index 0, 1, 2 = primes 2, 3, 5.
float radical_inverse(int base_index, uint64_t index) {
    const uint64_t divisor = base_index == 0 ? 2u : (base_index == 1 ? 3u : 5u);
    const float recip = 1.f / (float) divisor;
    uint64_t value = 0; float factor = 1.f;
    while (index != 0) {
        uint64_t next = index / divisor;
        factor = factor * recip;
        value = (value - next) * divisor + index;
        index = next;
    }
    float one_minus_eps; uint32_t u = 0x3f7fffffu; memcpy(&one_minus_eps, &u, 4);
    return std::min(one_minus_eps, (float) value * factor);
}

static float lerp(float a, float b, float t) { return fmadd(b, t, fmadd(-a, t, a)); }   // drjit/array_router.h:332

// bound_exit_pupil + compute_exit_pupil_bounds (realisticlens.cpp:1685-1809). The reference walks the 2^20 rays of a
// segment in order and skips those whose rear-plane point already lies in the box; the box is the min / max over the
// transmitted points either way, so the rays are simply dealt over threads here.
void Camera::compute_exit_pupil_bounds() {
    const size_t num_segments = 64, rays_per_segment = 1024 * 1024;
    const float R = ifs.front().elem_radius * 1.5f, rz = ifs.front().z;
    for (size_t seg = 0; seg < num_segments; ++seg) {
        const float r0 = seg * film_diagonal / num_segments, r1 = (seg + 1) * film_diagonal / num_segments;
        float mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY; long transmitted = 0;
        #pragma omp parallel for reduction(min : mnx, mny) reduction(max : mxx, mxy) reduction(+ : transmitted) schedule(static)
        for (long i = 0; i < (long) rays_per_segment; ++i) {
            V3 p_film { lerp(r0, r1, ((size_t) i + 0.5f) / rays_per_segment), 0.f, 0.f };
            V3 p_rear { lerp(-R, R, radical_inverse(0, (uint64_t) i)), lerp(-R, R, radical_inverse(1, (uint64_t) i)), rz };
            Ray ray { p_film, normalize(p_rear - p_film) }, ro;
            if (!trace_from_film(ray, ro)) continue;
            mnx = std::min(mnx, p_rear.x); mny = std::min(mny, p_rear.y); mxx = std::max(mxx, p_rear.x); mxy = std::max(mxy, p_rear.y);
            transmitted++;
        }
        if (transmitted == 0) { mnx = mny = -R; mxx = mxy = R; }
        const float spacing = 4 * R * std::sqrt(2.f / rays_per_segment);
        pupil_bounds[seg][0] = mnx - spacing; pupil_bounds[seg][1] = mny - spacing;
        pupil_bounds[seg][2] = mxx + spacing; pupil_bounds[seg][3] = mxy + spacing;
    }
}

// sample_exit_pupil (realisticlens.cpp:1812-1845)
V3 Camera::sample_exit_pupil(V3 film_p, V2 ap, float &bounds_area) const {
    float r_film = std::sqrt(sqr(film_p.x) + sqr(film_p.y));
    int idx = (int) std::floor(r_film / film_diagonal * 64);
    idx = std::min(std::max(idx, 0), 63);
    const float *b = pupil_bounds[idx];
    bounds_area = (b[2] - b[0]) * (b[3] - b[1]);
    float px = lerp(b[0], b[2], ap.x), py = lerp(b[1], b[3], ap.y);
    float sin_t = r_film > 0.f ? film_p.y * rcp(r_film) : 0.f, cos_t = r_film > 0.f ? film_p.x * rcp(r_film) : 1.f;
    return { cos_t * px - sin_t * py, sin_t * px + cos_t * py, rear_z };
}

// realisticlens.cpp:2209-2292 (rear-element sampling :1870-1878, or the exit-pupil LUT :1812-1845)
float Camera::sample_ray(V2 pos, V2 ap, Ray &out) const {
    V3 film_p = xform_point(sample_to_film, { pos.x, pos.y, 0.f });   // w == 1: the /w is exact
    float bounds_area = 0.f;
    V3 aperture_p;
    if (pupil_mode) aperture_p = sample_exit_pupil(film_p, ap, bounds_area);
    else {
        V2 tmp = square_to_uniform_disk_concentric(ap);
        aperture_p = { rear_radius * tmp.x, rear_radius * tmp.y, rear_z };
    }
    Ray ray { film_p, normalize(aperture_p - film_p) };
    Ray ro; bool active = trace_from_film(ray, ro);
    V3 d_local = ro.d;
    float weight = active ? 1.f : 0.f;
    ro.o.z += film_z;
    ro.o = xform_point(to_world, ro.o);
    ro.d = xform_vector(to_world, d_local);
    float inv_z = rcp(d_local.z);
    float near_t = near_clip * inv_z, far_t = far_clip * inv_z;
    ro.o = { fmadd(ro.d.x, near_t, ro.o.x), fmadd(ro.d.y, near_t, ro.o.y), fmadd(ro.d.z, near_t, ro.o.z) };
    ro.maxt = far_t - near_t;
    float ct = d_local.z, cos4t = sqr(sqr(ct));
    if (pupil_mode) weight *= shutter_open_time * bounds_area * rcp(sqr(rear_z));   // :2284-2286
    else weight *= cos4t;
    out = ro;
    return weight;
}

}  // namespace orc

// ================================================================================ C entry points
namespace orc {
Interface make_sphere(float, float, float, float, float);
Interface make_plane(int, float, float, float, float);
}
using namespace orc;

extern "C" {

void orc_sincos(int n, const float *x, float *s, float *c) { for (int i = 0; i < n; ++i) sincos(x[i], s[i], c[i]); }
void orc_concentric_disk(int n, const float *uv, float *out2) {
    for (int i = 0; i < n; ++i) { V2 p = square_to_uniform_disk_concentric({ uv[2 * i], uv[2 * i + 1] }); out2[2 * i] = p.x; out2[2 * i + 1] = p.y; }
}
void orc_cosine_hemisphere(int n, const float *uv, float *out3) {
    for (int i = 0; i < n; ++i) { V3 p = square_to_cosine_hemisphere({ uv[2 * i], uv[2 * i + 1] }); out3[3 * i] = p.x; out3[3 * i + 1] = p.y; out3[3 * i + 2] = p.z; }
}
void orc_fresnel(int n, const float *cos_i, const float *eta, float *out4) {
    for (int i = 0; i < n; ++i) fresnel(cos_i[i], eta[i], out4[4 * i], out4[4 * i + 1], out4[4 * i + 2], out4[4 * i + 3]);
}
void orc_refract(int n, const float *wi, const float *m, const float *ct, const float *eti, float *out3) {
    for (int i = 0; i < n; ++i) {
        V3 r = refract({ wi[3 * i], wi[3 * i + 1], wi[3 * i + 2] }, { m[3 * i], m[3 * i + 1], m[3 * i + 2] }, ct[i], eti[i]);
        out3[3 * i] = r.x; out3[3 * i + 1] = r.y; out3[3 * i + 2] = r.z;
    }
}
void orc_solve_quadratic(int n, const float *abc, float *out3) {
    for (int i = 0; i < n; ++i) {
        float x0, x1; bool v = solve_quadratic(abc[3 * i], abc[3 * i + 1], abc[3 * i + 2], x0, x1);
        out3[3 * i] = v ? 1.f : 0.f; out3[3 * i + 1] = x0; out3[3 * i + 2] = x1;
    }
}
float orc_material_ior(const float *mat7, float wavelength_um) { return material_ior(mat7, wavelength_um); }

// film: width,height, crop (w,h,ox,oy), diagonal in mm (film.cpp:46-48 stores 0.001*diag)
static Camera *new_camera(int width, int height, const int *crop4, float diagonal_mm, const float *to_world16,
                          float near_clip, float far_clip) {
    Camera *c = new Camera();
    setup_film(*c, width, height, crop4[0], crop4[1], crop4[2], crop4[3], 0.001f * diagonal_mm);
    std::memcpy(c->to_world, to_world16, 64);
    c->near_clip = near_clip; c->far_clip = far_clip;
    return c;
}
// prescription tables (front -> rear, mm), materials (n+1) x 7
void *orc_cam_from_prescription(int n, const int *kinds, const float *thick, const float *curv, const float *elemr,
                                const float *kappa, const float *ai7, const float *mats7, float scale,
                                float stop_ratio, int fine_focus, float object_distance, int width, int height,
                                const int *crop4, float diagonal_mm, const float *to_world16, float near_clip,
                                float far_clip) {
    Camera *c = new_camera(width, height, crop4, diagonal_mm, to_world16, near_clip, far_clip);
    build_from_prescription(*c, n, kinds, thick, curv, elemr, kappa, ai7, mats7, scale, stop_ratio);
    finish_camera(*c, fine_focus != 0, object_distance);
    return c;
}
// ready-made interfaces, rear -> front, metres: rows of 16 floats
// [kind, z, elem_radius, curv_radius | (c, K, r_scale), ai0..6 at [6..12], ior_left, ior_right, -]
void *orc_cam_from_interfaces(int n, const float *rows16, float film_z, int fine_focus, float object_distance,
                              int width, int height, const int *crop4, float diagonal_mm, const float *to_world16,
                              float near_clip, float far_clip) {
    Camera *c = new_camera(width, height, crop4, diagonal_mm, to_world16, near_clip, far_clip);
    for (int i = 0; i < n; ++i) {
        const float *r = rows16 + 16 * i; int kind = (int) r[0];
        Interface f;
        if (kind == SPHERE) f = make_sphere(r[3], r[2], r[1], r[13], r[14]);
        else if (kind == ASPHERE) {
            f.kind = ASPHERE; f.z = r[1]; f.elem_radius = r[2]; f.c = r[3]; f.K = r[4]; f.r_scale = r[5];
            for (int k = 0; k < 7; ++k) f.ai[k] = r[6 + k];
            f.ior_left = r[13]; f.ior_right = r[14];
        } else f = make_plane(kind, r[2], r[1], r[13], r[14]);
        c->ifs.push_back(f);
    }
    c->film_z = film_z;
    finish_camera(*c, fine_focus != 0, object_distance);
    return c;
}
void orc_cam_destroy(void *h) { delete (Camera *) h; }
int orc_cam_num_interfaces(void *h) { return (int) ((Camera *) h)->ifs.size(); }
float orc_cam_focus_delta(void *h) { return ((Camera *) h)->focus_delta; }
void orc_cam_sample_to_film(void *h, float *out16) { std::memcpy(out16, ((Camera *) h)->sample_to_film, 64); }
void orc_cam_film_to_sample(void *h, float *out16) { std::memcpy(out16, ((Camera *) h)->film_to_sample, 64); }
// same 16-float row as oracle/_ref's ref_cam_interface
void orc_cam_interface(void *h, int i, float *o) {
    const Interface &f = ((Camera *) h)->ifs[i];
    for (int k = 0; k < 16; ++k) o[k] = 0.f;
    o[0] = (float) f.kind; o[1] = f.z; o[2] = f.elem_radius;
    if (f.kind == SPHERE) o[3] = f.curv_radius;
    if (f.kind == ASPHERE) { o[3] = f.c; o[4] = f.K; o[5] = f.r_scale; for (int k = 0; k < 7; ++k) o[6 + k] = f.ai[k]; }
    o[13] = f.ior_left; o[14] = f.ior_right;
}
float orc_radical_inverse(int base_index, uint64_t index) { return radical_inverse(base_index, index); }
void orc_cam_set_exit_pupil(void *h, int enable, float shutter_open_time) {
    Camera *c = (Camera *) h;
    c->shutter_open_time = shutter_open_time;
    if (enable) c->compute_exit_pupil_bounds();
    c->pupil_mode = enable != 0;
}
void orc_cam_get_exit_pupil(void *h, float *out256) { memcpy(out256, ((Camera *) h)->pupil_bounds, 64 * 4 * sizeof(float)); }
// out10 = o[3] d[3] maxt weight[3]
void orc_cam_sample_ray(void *h, int n, const float *pos2, const float *ap2, float *out10) {
    const Camera *c = (const Camera *) h;
    for (int i = 0; i < n; ++i) {
        Ray r; float w = c->sample_ray({ pos2[2 * i], pos2[2 * i + 1] }, { ap2[2 * i], ap2[2 * i + 1] }, r);
        float *o = out10 + 10 * i;
        o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
        o[6] = r.maxt; o[7] = o[8] = o[9] = w;
    }
}
// in o[3] d[3]; out o[3] d[3] active
void orc_cam_trace(void *h, int from_world, int n, const float *od6, float *out7) {
    const Camera *c = (const Camera *) h;
    for (int i = 0; i < n; ++i) {
        const float *r = od6 + 6 * i;
        Ray in { { r[0], r[1], r[2] }, { r[3], r[4], r[5] } }, ro;
        bool a = from_world ? c->trace_from_world(in, ro) : c->trace_from_film(in, ro);
        float *o = out7 + 7 * i;
        o[0] = ro.o.x; o[1] = ro.o.y; o[2] = ro.o.z; o[3] = ro.d.x; o[4] = ro.d.y; o[5] = ro.d.z; o[6] = a ? 1.f : 0.f;
    }
}
}  // extern "C"
