// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). The lens-design spot path:
//   * Camera::sample_direction — float32 restatement of RealisticLensCamera::sample_direction
//     (src/sensors/realisticlens.cpp:2301-2394), pinned bit-for-bit against oracle/_ref (tests/golden/lens.npz).
//   * orc_spot_render — the ptracer connection of a point emitter to that sensor: sample_direction +
//     ImageBlock::put(normalize=true, coalesce=false) with sample_scale 1/spp
//     (src/integrators/ptracer.cpp connect_sensor; src/render/integrator.cpp:671-676,706-717).
//   * orc_spot_grad — the derivative checker for the CUDA adjoint. Deliberately a DIFFERENT method from the
//     product's hand-derived reverse sweep: forward-mode dual numbers over float64 pushed through a generic
//     re-statement of the same lens classes, seeded on the selected (interface, column) entries of the table.
//     The reference has no derivative of this code (its lens tables are constants), so this checker plus central
//     finite differences of orc_spot_render-in-double (tests) are what pin the gradient.
#include "oracle.h"
#include "oracle_film.h"
#include <cstring>
#include <algorithm>

namespace orc {

static V3 xform_affine_diag(const float *m, V3 p) {   // Transform4f::transform_affine, column fmadd chain
    V3 r { m[3], m[7], m[11] };
    r = { fmadd(m[0], p.x, r.x), fmadd(m[4], p.x, r.y), fmadd(m[8], p.x, r.z) };
    r = { fmadd(m[1], p.y, r.x), fmadd(m[5], p.y, r.y), fmadd(m[9], p.y, r.z) };
    r = { fmadd(m[2], p.z, r.x), fmadd(m[6], p.z, r.y), fmadd(m[10], p.z, r.z) };
    return r;
}

// realisticlens.cpp:2301-2394, camera space (to_world = identity in the tests that call this)
// origin_offset > 0: the product's well-conditioned variant (same ray, started near the lens; lrt_lens.cuh)
bool camera_sample_direction(const Camera &c, V3 ref_p, V2 sample, float *out11, float origin_offset = 0.f) {
    for (int i = 0; i < 11; ++i) out11[i] = 0.f;
    bool active = ref_p.z >= c.near_clip && ref_p.z <= c.far_clip;
    if (!active) return false;
    const Interface &front = c.ifs.back();
    float front_radius = 1.0f * front.elem_radius, front_z = front.z;
    V2 tmp = square_to_uniform_disk_concentric(sample);
    V3 aperture_p { tmp.x * front_radius, tmp.y * front_radius, front_z };
    V3 dir = ref_p - aperture_p;
    float dist = norm(dir), inv_dist = rcp(dist);
    dir = dir * inv_dist;
    Ray world_ray { ref_p, -dir }, ray_out;
    if (origin_offset > 0.f)
        world_ray.o = { fmadd(dir.x, origin_offset, aperture_p.x), fmadd(dir.y, origin_offset, aperture_p.y), fmadd(dir.z, origin_offset, aperture_p.z) };
    bool valid = c.trace_from_world(world_ray, ray_out);
    if (!valid) return false;
    float t = -ray_out.o.z / ray_out.d.z;
    V3 pf { fmadd(ray_out.d.x, t, ray_out.o.x), fmadd(ray_out.d.y, t, ray_out.o.y), fmadd(ray_out.d.z, t, ray_out.o.z) };
    V3 scr = xform_affine_diag(c.film_to_sample, pf);
    valid = valid && scr.x >= 0.f && scr.y >= 0.f && scr.x <= 1.f && scr.y <= 1.f;
    float ct_film = -ray_out.d.z, inv_ct = rcp(ct_film);
    float value = valid ? c.normalization * sqr(sqr(inv_ct)) * sqr(c.rear_z) : 0.f;
    if (!valid) return false;
    out11[0] = scr.x * c.res_x; out11[1] = scr.y * c.res_y;
    out11[2] = aperture_p.x; out11[3] = aperture_p.y; out11[4] = aperture_p.z;
    V3 d = (aperture_p - ref_p) * inv_dist;
    out11[5] = d.x; out11[6] = d.y; out11[7] = d.z; out11[8] = dist;
    float aperture_pdf = rcp(Pi * sqr(front_radius));
    float ct_ref = dir.z;
    out11[9] = aperture_pdf * dist * dist * rcp(ct_ref);
    out11[10] = value * inv_dist * inv_dist * ct_ref;
    return true;
}

// ================================================================================ generic-scalar restatement
// Dual numbers: value + up to kMaxSeeds partial derivatives (forward mode).
constexpr int kMaxSeeds = 64;
static thread_local int g_nseeds = 0;
struct Dual {
    double v; double d[kMaxSeeds];
    Dual(double x = 0.0) : v(x) { for (int i = 0; i < g_nseeds; ++i) d[i] = 0.0; }
};
static Dual operator+(const Dual &a, const Dual &b) { Dual r(a.v + b.v); for (int i = 0; i < g_nseeds; ++i) r.d[i] = a.d[i] + b.d[i]; return r; }
static Dual operator-(const Dual &a, const Dual &b) { Dual r(a.v - b.v); for (int i = 0; i < g_nseeds; ++i) r.d[i] = a.d[i] - b.d[i]; return r; }
static Dual operator-(const Dual &a) { Dual r(-a.v); for (int i = 0; i < g_nseeds; ++i) r.d[i] = -a.d[i]; return r; }
static Dual operator*(const Dual &a, const Dual &b) { Dual r(a.v * b.v); for (int i = 0; i < g_nseeds; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
static Dual operator/(const Dual &a, const Dual &b) {
    Dual r(a.v / b.v); double ib = 1.0 / b.v;
    for (int i = 0; i < g_nseeds; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * ib;
    return r;
}
static Dual dsqrt(const Dual &a) { Dual r(std::sqrt(a.v)); double h = 0.5 / r.v; for (int i = 0; i < g_nseeds; ++i) r.d[i] = a.d[i] * h; return r; }
static Dual dabs(const Dual &a) { return a.v < 0 ? -a : a; }

struct DV3 { Dual x, y, z; };
static DV3 operator+(const DV3 &a, const DV3 &b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
static DV3 operator-(const DV3 &a, const DV3 &b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
static DV3 operator*(const DV3 &a, const Dual &s) { return { a.x * s, a.y * s, a.z * s }; }
static Dual ddot(const DV3 &a, const DV3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static DV3 dnormalize(const DV3 &a) { return a * (Dual(1.0) / dsqrt(ddot(a, a))); }

struct DIf {   // one interface with differentiable parameters
    int kind; double elem_radius;
    Dual z, R, c, K, ai[7], ior_l, ior_r;
};
struct DRay { DV3 o, d; };

static Dual asph_S(const DIf &f, const Dual &u) {         // realisticlens.cpp:680-704 in unitless radius
    Dual P(0.0);
    for (int i = 6; i >= 0; --i) P = P * u + f.ai[i];
    Dual w = dsqrt(Dual(1.0) - (Dual(1.0) + f.K) * f.c * f.c * u);
    return P * u * u + f.c * u / (Dual(1.0) + w);
}
static Dual asph_dS_du(const DIf &f, const Dual &u) {
    Dual P(0.0), Pp(0.0);
    for (int i = 6; i >= 0; --i) { Pp = Pp * u + P; P = P * u + f.ai[i]; }
    Dual w = dsqrt(Dual(1.0) - (Dual(1.0) + f.K) * f.c * f.c * u);
    return Pp * u * u + Dual(2.0) * u * P + f.c / (Dual(2.0) * w);
}

// LensInterface::compute_interaction (:269-297) over Dual; masks decided on the values
static bool d_interact(const DIf &f, DRay &ray) {
    DV3 p, n; Dual t;
    if (f.kind == SPHERE) {   // :359-402
        DV3 center { Dual(0.0), Dual(0.0), f.z + f.R };
        DV3 pcl = center - ray.o;
        Dual proj = ddot(pcl, ray.d);
        DV3 perp = pcl - ray.d * proj;
        Dual disc = f.R * f.R - ddot(perp, perp);
        if (!(disc.v >= 0)) return false;
        Dual sq = dsqrt(disc), near_t = proj - sq, far_t = proj + sq;
        if (!(far_t.v >= 0)) return false;
        bool convex = f.R.v > 0;
        if (convex != (ray.d.z.v < 0)) { if (!(near_t.v >= 0)) return false; t = near_t; } else t = far_t;
        p = ray.o + ray.d * t;
        n = dnormalize(p - center);
        if (!convex) n = n * Dual(-1.0);
    } else if (f.kind == ASPHERE) {   // :574-645 — Newton on F(t) = p.z - sag(p.x, p.y), run to float64 convergence
        double Re = f.elem_radius;
        t = (f.z - ray.o.z) / ray.d.z;
        // start from the value-level conic/plane guess refined by many Newton steps (derivatives ride along)
        bool ok = false;
        for (int it = 0; it < 60; ++it) {
            p = ray.o + ray.d * t;
            Dual u = (p.x * p.x + p.y * p.y) / Dual(Re * Re);
            if (!((Dual(1.0) - (Dual(1.0) + f.K) * f.c * f.c * u).v > 0)) return false;
            Dual sag = f.z - Dual(Re) * asph_S(f, u);
            Dual F = p.z - sag;
            Dual k = Dual(-2.0 / Re) * asph_dS_du(f, u);          // sag_x = k x
            Dual dF = ray.d.z - k * (p.x * ray.d.x + p.y * ray.d.y);
            Dual step = F / dF;
            t = t - step;
            if (std::fabs(step.v) < 1e-15 && it > 2) { ok = true; break; }
        }
        if (!ok || !(t.v >= 0)) return false;
        p = ray.o + ray.d * t;
        Dual u = (p.x * p.x + p.y * p.y) / Dual(Re * Re);
        Dual k = Dual(-2.0 / Re) * asph_dS_du(f, u);
        n = dnormalize(DV3 { k * p.x, k * p.y, Dual(-1.0) });
    } else {   // :444-517
        if (!(std::fabs(ray.d.z.v) >= Epsilon)) return false;
        t = (f.z - ray.o.z) / ray.d.z;
        if (!(t.v >= 0)) return false;
        p = ray.o + ray.d * t;
        n = DV3 { Dual(0.0), Dual(0.0), Dual(-1.0) };
    }
    if (!(p.x.v * p.x.v + p.y.v * p.y.v < f.elem_radius * f.elem_radius)) return false;
    if (f.kind == APERTURE) { ray.o = p; return true; }
    Dual ci = -ddot(ray.d, n);
    Dual eta = f.ior_r / f.ior_l;
    bool outside = ci.v >= 0;
    Dual eti = outside ? Dual(1.0) / eta : eta;
    Dual q = Dual(1.0) - (Dual(1.0) - ci * ci) * eti * eti;
    if (!(q.v > 0)) return false;                                   // total internal reflection
    Dual ct = dsqrt(q); if (outside) ct = -ct;
    Dual g = ci * eti + ct;
    ray.d = n * g + ray.d * eti;
    ray.o = p;
    return true;
}

struct DFilter {
    int kind; double radius; double coeff[10];
    Dual eval(const Dual &x) const {
        if (kind == 1) { Dual r = Dual(1.0) - dabs(x) / Dual(radius); return r.v > 0 ? r : Dual(0.0); }
        Dual x2 = x * x, P(0.0);
        for (int i = 9; i >= 0; --i) P = P * x2 + Dual(coeff[i]);
        return P.v > 0 ? P : Dual(0.0);
    }
};

}  // namespace orc

using namespace orc;

extern "C" {

void orc_cam_sample_direction(void *h, int n, const float *p3, const float *s2, float *out11) {
    const Camera &c = *(Camera *) h;
    for (int i = 0; i < n; ++i)
        camera_sample_direction(c, { p3[3 * i], p3[3 * i + 1], p3[3 * i + 2] }, { s2[2 * i], s2[2 * i + 1] }, out11 + 11 * i);
}

// Jacobian of one lens trace in float64 forward mode: seeds = the 6 ray inputs (o, d; d is NOT renormalised)
// followed by the n_sel selected table entries. jac[6][6 + n_sel] row-major; returns the `active` mask.
static void load_ifs(int n_if, const float *rows16, std::vector<DIf> &ifs) {
    ifs.resize(n_if);
    for (int i = 0; i < n_if; ++i) {
        const float *r = rows16 + 16 * i; DIf &f = ifs[i];
        f.kind = (int) r[0]; f.elem_radius = r[2]; f.z = Dual(r[1]); f.R = Dual(r[3]); f.c = Dual(r[3]); f.K = Dual(r[4]);
        for (int k = 0; k < 7; ++k) f.ai[k] = Dual(r[6 + k]);
        f.ior_l = Dual(r[13]); f.ior_r = Dual(f.kind == APERTURE ? r[13] : r[14]);
    }
}
static Dual *sel_target(DIf &f, int col) {
    return col == 1 ? &f.z : col == 3 ? (f.kind == ASPHERE ? &f.c : &f.R) : col == 4 ? &f.K
           : (col >= 6 && col <= 12) ? &f.ai[col - 6] : col == 13 ? &f.ior_l : &f.ior_r;
}
int orc_trace_jac(int n_if, const float *rows16, int from_world, const float *od6, int n_sel, const int *sel, double *out6, double *jac) {
    g_nseeds = 6 + n_sel;
    std::vector<DIf> ifs; load_ifs(n_if, rows16, ifs);
    for (int s = 0; s < n_sel; ++s) sel_target(ifs[sel[2 * s]], sel[2 * s + 1])->d[6 + s] = 1.0;
    Dual in[6];
    for (int k = 0; k < 6; ++k) { in[k] = Dual(od6[k]); in[k].d[k] = 1.0; }
    DRay ray { { in[0], in[1], in[2] }, { in[3], in[4], in[5] } };
    bool ok = true;
    for (int q = 0; q < n_if && ok; ++q) ok = d_interact(ifs[from_world ? n_if - 1 - q : q], ray);
    if (!ok) return 0;
    const Dual *o[6] = { &ray.o.x, &ray.o.y, &ray.o.z, &ray.d.x, &ray.d.y, &ray.d.z };
    for (int r = 0; r < 6; ++r) { out6[r] = o[r]->v; for (int c = 0; c < g_nseeds; ++c) jac[r * g_nseeds + c] = o[r]->d[c]; }
    return 1;
}

// ptracer-style spot image in float64 accumulators: storage [size_y, size_x, 4]; returns the number of valid samples
long long orc_spot_render(void *h, int size_x, int size_y, int off_x, int off_y, int filter_kind, float filter_param,
                          uint32_t seed, int spp, const float *src3, const float *intensity3, float origin_offset, double *storage);

// d/d(rows16[sel]) of  Phi = sum_samples sum_taps G[tap] . contribution  (masks frozen), forward-mode in float64.
// rows16: [n_if,16] interface table (rear -> front); sel: n_sel pairs (interface, column); grad_out[n_sel].
// Also returns Phi itself (value) so that tests can finite-difference it.
double orc_spot_grad(int n_if, const float *rows16, const float *f2s16, float normalization, float near_clip, float far_clip,
                     int size_x, int size_y, int off_x, int off_y, int filter_kind, float filter_param, uint32_t seed, int spp,
                     const float *src3, const float *intensity3, float origin_offset, const float *G /* [sy,sx,4] */, int n_sel, const int *sel,
                     const double *delta /* optional [n_sel] perturbation added to the selected entries */, double *grad_out) {
    g_nseeds = std::min(n_sel, kMaxSeeds);
    std::vector<DIf> ifs(n_if);
    for (int i = 0; i < n_if; ++i) {
        const float *r = rows16 + 16 * i; DIf &f = ifs[i];
        f.kind = (int) r[0]; f.elem_radius = r[2]; f.z = Dual(r[1]); f.R = Dual(r[3]); f.c = Dual(r[3]); f.K = Dual(r[4]);
        for (int k = 0; k < 7; ++k) f.ai[k] = Dual(r[6 + k]);
        f.ior_l = Dual(r[13]); f.ior_r = Dual(f.kind == APERTURE ? r[13] : r[14]);
    }
    for (int s = 0; s < g_nseeds; ++s) {
        DIf &f = ifs[sel[2 * s]]; int col = sel[2 * s + 1];
        Dual *tgt = col == 1 ? &f.z : col == 3 ? (f.kind == ASPHERE ? &f.c : &f.R) : col == 4 ? &f.K
                    : (col >= 6 && col <= 12) ? &f.ai[col - 6] : col == 13 ? &f.ior_l : &f.ior_r;
        tgt->d[s] = 1.0;
        if (delta) tgt->v += delta[s];
    }
    DFilter flt; flt.kind = filter_kind;
    if (filter_kind == 1) flt.radius = filter_param > 0 ? filter_param : 1.0;
    else {
        double stddev = filter_param > 0 ? filter_param : 0.5; flt.radius = 4 * stddev;
        const double c[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                               -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
        double scale = 1; for (int i = 0; i < 10; ++i) { flt.coeff[i] = c[i] * scale; scale /= stddev * stddev; }
        double x2 = flt.radius * flt.radius, P = 0; for (int i = 9; i >= 0; --i) P = P * x2 + flt.coeff[i];
        flt.coeff[0] -= P;
    }
    const int count_max = (int) std::ceil(2 * flt.radius);
    const long long n = (long long) size_x * size_y * spp;
    const double scale = 1.0 / spp;
    Dual phi(0.0);
    for (long long idx = 0; idx < n; ++idx) {
        PCG32 rng; sampler_seed_jit(rng, seed, (uint32_t) idx);
        float sa = rng.next_float32(), sb = rng.next_float32();
        if (!(src3[2] >= near_clip && src3[2] <= far_clip)) continue;
        V2 dsk = square_to_uniform_disk_concentric({ sa, sb });
        const DIf &front = ifs[n_if - 1];
        DV3 ap { Dual(dsk.x * front.elem_radius), Dual(dsk.y * front.elem_radius), front.z };
        DV3 src { Dual(src3[0]), Dual(src3[1]), Dual(src3[2]) };
        DV3 v = src - ap;
        Dual dist = dsqrt(ddot(v, v));
        DV3 dir = v * (Dual(1.0) / dist);
        DRay ray { origin_offset > 0.f ? ap + dir * Dual(origin_offset) : src, dir * Dual(-1.0) };
        bool ok = true;
        for (int i = n_if - 1; i >= 0 && ok; --i) ok = d_interact(ifs[i], ray);
        if (!ok) continue;
        Dual t = -ray.o.z / ray.d.z;
        DV3 pf = ray.o + ray.d * t;
        Dual sx = Dual(f2s16[0]) * pf.x + Dual(f2s16[3]), sy = Dual(f2s16[5]) * pf.y + Dual(f2s16[7]);
        if (!(sx.v >= 0 && sy.v >= 0 && sx.v <= 1 && sy.v <= 1)) continue;
        Dual ct = -ray.d.z;
        Dual value = Dual(normalization) / (ct * ct * ct * ct) * ifs[0].z * ifs[0].z;
        Dual wgt = value / (dist * dist) * dir.z;
        // ImageBlock::put, non-coalesced + normalize (imageblock.cpp:264-412); block without border
        Dual px = sx * Dual(size_x) + Dual(off_x), py = sy * Dual(size_y) + Dual(off_y);
        Dual pfx = px - Dual(off_x + 0.5), pfy = py - Dual(off_y + 0.5);
        int cx = (int) std::ceil(pfx.v - flt.radius), cy = (int) std::ceil(pfy.v - flt.radius);
        int x0 = std::max(cx, 0), y0 = std::max(cy, 0);
        int x1 = std::min((int) std::floor(pfx.v + flt.radius), size_x - 1), y1 = std::min((int) std::floor(pfy.v + flt.radius), size_y - 1);
        if (x0 > x1 || y0 > y1) continue;
        Dual a(0.0), b(0.0);
        for (int i = 0; i < count_max; ++i) { a = a + flt.eval(Dual(cx + i) - pfx); b = b + flt.eval(Dual(cy + i) - pfy); }
        if (a.v == 0 || b.v == 0) continue;
        int nx = std::min(x1 - x0 + 1, count_max), ny = std::min(y1 - y0 + 1, count_max);
        for (int iy = 0; iy < ny; ++iy)
            for (int ix = 0; ix < nx; ++ix) {
                Dual w = flt.eval(Dual(x0 + ix) - pfx) * flt.eval(Dual(y0 + iy) - pfy) / (a * b);
                const float *g = G + ((size_t) (y0 + iy) * size_x + (x0 + ix)) * 4;
                Dual val = (Dual(g[0] * intensity3[0] + g[1] * intensity3[1] + g[2] * intensity3[2]) * wgt + Dual(g[3])) * Dual(scale);
                phi = phi + val * w;
            }
    }
    for (int s = 0; s < g_nseeds; ++s) grad_out[s] = phi.d[s];
    return phi.v;
}

}  // extern "C"

// float32 spot image (the checker of lrt_spot_render): pinned sample_direction + the oracle's ImageBlock::put
extern "C" long long orc_spot_render(void *h, int size_x, int size_y, int off_x, int off_y, int filter_kind, float filter_param,
                                     uint32_t seed, int spp, const float *src3, const float *intensity3, float origin_offset, double *storage) {
    const Camera &c = *(Camera *) h;
    orc::Filter flt; flt.init(filter_kind, filter_param);
    const long long n = (long long) size_x * size_y * spp;
    const float scale = 1.f / (float) spp;
    long long valid = 0;
    for (long long idx = 0; idx < n; ++idx) {
        PCG32 rng; sampler_seed_jit(rng, seed, (uint32_t) idx);
        float sa = rng.next_float32(), sb = rng.next_float32();
        float o[11];
        if (!camera_sample_direction(c, { src3[0], src3[1], src3[2] }, { sa, sb }, o, origin_offset)) continue;
        valid++;
        float wv = o[10] * scale;
        float vals[4] = { intensity3[0] * wv, intensity3[1] * wv, intensity3[2] * wv, scale };
        block_put(flt, size_x, size_y, off_x, off_y, 0, 4, false, true, false, o[0] + (float) off_x, o[1] + (float) off_y, vals, storage);
    }
    return valid;
}
