// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). Integrator-side restatement:
//   SamplingIntegrator::render / render_sample, JIT branch   src/render/integrator.cpp:247-330,419-491
//   PathIntegrator::sample                                   src/integrators/path.cpp:95-305
//   Scene::sample_emitter_direction / pdf_emitter_direction  src/render/scene.cpp:227-366
//   AreaLight                                                src/emitters/area.cpp:83-211
//   Shape::sample_direction / pdf_direction                  src/render/shape.cpp:358-388
//   Rectangle::sample_position                               src/shapes/rectangle.cpp:166-185
//   Mesh Moeller-Trumbore + compute_surface_interaction      include/mitsuba/render/mesh.h:432-452,
//                                                            src/render/mesh.cpp:1376-1500
//   Interaction::spawn_ray / spawn_ray_to / offset_p         include/mitsuba/render/interaction.h:133-162
//   SmoothDiffuse / SmoothConductor / SmoothDielectric       src/bsdfs/{diffuse,conductor,dielectric}.cpp
//   TwoSidedBRDF                                             src/bsdfs/twosided.cpp
// Geometry model of this build (product and oracle alike): every shape is a triangle list (rectangles and
// cubes are tessellated by the host front end); area emitters sit on rectangles and are sampled through
// the rectangle's own parameterisation. Closest hits come from a plain binary BVH here (the product uses
// its own 8-wide quantised BVH) — the closest hit is backend-independent up to exact ties.
// This part is "parity unpinned" against a build of the full reference (see oracle.h / DESIGN.md).
#include "oracle.h"
#include <algorithm>
#include <cstring>
#include <cstdio>

namespace orc {

struct Tri { V3 p0, p1, p2; int shape; };
struct Bsdf { int type; bool twosided; V3 refl, eta, k, trans; float eta_s; };   // 0 diffuse 1 conductor 2 dielectric
struct Emitter { V3 radiance; int shape; V3 c, ex, ey, n; float inv_area; };

struct BNode { V3 lo, hi; int left, right, first, count; };

struct Scene {
    std::vector<Tri> tris;
    std::vector<int> shape_bsdf, shape_emitter;
    std::vector<Bsdf> bsdfs;
    std::vector<Emitter> emitters;
    std::vector<BNode> nodes;
    std::vector<int> order;
    void build();
    int build_rec(int first, int count);
};

static V3 vmin(V3 a, V3 b) { return { std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z) }; }
static V3 vmax(V3 a, V3 b) { return { std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z) }; }
static float comp(V3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

int Scene::build_rec(int first, int count) {
    BNode nd; nd.lo = { Infinity, Infinity, Infinity }; nd.hi = { -Infinity, -Infinity, -Infinity };
    V3 clo = nd.lo, chi = nd.hi;
    for (int i = first; i < first + count; ++i) {
        const Tri &t = tris[order[i]];
        nd.lo = vmin(nd.lo, vmin(t.p0, vmin(t.p1, t.p2))); nd.hi = vmax(nd.hi, vmax(t.p0, vmax(t.p1, t.p2)));
        V3 c = (t.p0 + t.p1 + t.p2) * (1.f / 3.f);
        clo = vmin(clo, c); chi = vmax(chi, c);
    }
    nd.left = nd.right = -1; nd.first = first; nd.count = count;
    int id = (int) nodes.size(); nodes.push_back(nd);
    if (count > 4) {
        V3 e = chi - clo; int axis = e.x > e.y ? (e.x > e.z ? 0 : 2) : (e.y > e.z ? 1 : 2);
        int mid = first + count / 2;
        std::nth_element(order.begin() + first, order.begin() + mid, order.begin() + first + count, [&](int a, int b) {
            const Tri &ta = tris[a], &tb = tris[b];
            return comp(ta.p0 + ta.p1 + ta.p2, axis) < comp(tb.p0 + tb.p1 + tb.p2, axis);
        });
        int l = build_rec(first, mid - first), r = build_rec(mid, first + count - mid);
        nodes[id].left = l; nodes[id].right = r; nodes[id].count = 0;
    }
    return id;
}
void Scene::build() {
    order.resize(tris.size());
    for (size_t i = 0; i < tris.size(); ++i) order[i] = (int) i;
    nodes.clear(); nodes.reserve(tris.size());
    if (!tris.empty()) build_rec(0, (int) tris.size());
}

// mesh.h:432-452
static bool moeller_trumbore(const Ray &ray, const Tri &tr, float &t, float &u, float &v) {
    V3 e1 = tr.p1 - tr.p0, e2 = tr.p2 - tr.p0;
    V3 pvec = cross(ray.d, e2);
    float inv_det = rcp(dot(e1, pvec));
    V3 tvec = ray.o - tr.p0;
    u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    v = dot(ray.d, qvec) * inv_det;
    active = active && v >= 0.f && u + v <= 1.f;
    t = dot(e2, qvec) * inv_det;
    active = active && t >= 0.f && t <= ray.maxt;
    return active;
}

static bool hit_box(const BNode &n, const Ray &r, V3 inv_d, float tmax) {
    float t0 = 0.f, t1 = tmax;
    for (int a = 0; a < 3; ++a) {
        float o = comp(r.o, a), id = comp(inv_d, a);
        float ta = (comp(n.lo, a) - o) * id, tb = (comp(n.hi, a) - o) * id;
        if (ta > tb) std::swap(ta, tb);
        // conservative slack so that the box test never culls a triangle hit by rounding
        t0 = std::fmax(t0, ta * (1.f - 4e-7f) - 1e-30f); t1 = std::fmin(t1, tb * (1.f + 4e-7f) + 1e-30f);
        if (std::isnan(ta) || std::isnan(tb)) { t0 = 0.f; }
    }
    return t0 <= t1;
}

struct Hit { float t = Infinity, u = 0, v = 0; int prim = -1; };

static Hit intersect(const Scene &s, const Ray &ray, bool any_hit) {
    Hit best; if (s.nodes.empty()) return best;
    Ray r = ray;
    V3 inv_d { 1.f / r.d.x, 1.f / r.d.y, 1.f / r.d.z };
    int stack[128], sp = 0; stack[sp++] = 0;
    while (sp) {
        const BNode &n = s.nodes[stack[--sp]];
        if (!hit_box(n, r, inv_d, r.maxt)) continue;
        if (n.left < 0) {
            for (int i = n.first; i < n.first + n.count; ++i) {
                float t, u, v;
                if (moeller_trumbore(r, s.tris[s.order[i]], t, u, v)) {
                    // ties: keep the lowest primitive index so that the result is order-independent
                    if (t < best.t || (t == best.t && s.order[i] < best.prim)) { best.t = t; best.u = u; best.v = v; best.prim = s.order[i]; }
                    if (any_hit) return best;
                    r.maxt = best.t;
                }
            }
        } else { stack[sp++] = n.left; stack[sp++] = n.right; }
    }
    return best;
}

// vector.h:116-136 coordinate_system (Duff et al.)
static void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = std::copysign(1.f, n.z);        // dr::sign
    float a = -rcp(sign + n.z), b = n.x * n.y * a;
    auto mulsign = [](float v, float sgn) { return std::signbit(sgn) ? -v : v; };
    s = { mulsign(sqr(n.x) * a, n.z) + 1, mulsign(b, n.z), n.z >= 0.f ? -n.x : n.x };
    t = { b, fmadd(n.y, n.y * a, sign), -n.y };
}

struct SI {
    float t = Infinity; V3 p, n, sh_s, sh_t, sh_n, wi; int shape = -1;
    bool valid() const { return t != Infinity; }
    V3 to_local(V3 v) const { return { dot(v, sh_s), dot(v, sh_t), dot(v, sh_n) }; }
    V3 to_world(V3 v) const {   // frame.h: fmadd(n, v.z, fmadd(t, v.y, s * v.x))
        return { fmadd(sh_n.x, v.z, fmadd(sh_t.x, v.y, sh_s.x * v.x)), fmadd(sh_n.y, v.z, fmadd(sh_t.y, v.y, sh_s.y * v.x)),
                 fmadd(sh_n.z, v.z, fmadd(sh_t.z, v.y, sh_s.z * v.x)) };
    }
};

// mesh.cpp:1376-1500 (no vertex normals / texcoords) + interaction.h:245-255,464-484
static SI surface_interaction(const Scene &s, const Ray &ray, const Hit &h) {
    SI si; if (h.prim < 0) return si;
    const Tri &tr = s.tris[h.prim];
    float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
    V3 pb = tr.p2 * b2;
    pb = { fmadd(tr.p1.x, b1, pb.x), fmadd(tr.p1.y, b1, pb.y), fmadd(tr.p1.z, b1, pb.z) };
    si.p = { fmadd(tr.p0.x, b0, pb.x), fmadd(tr.p0.y, b0, pb.y), fmadd(tr.p0.z, b0, pb.z) };
    si.t = h.t;
    si.n = normalize(cross(tr.p1 - tr.p0, tr.p2 - tr.p0));
    V3 dp_du, dp_dv; coordinate_system(si.n, dp_du, dp_dv);
    si.sh_n = si.n;
    float d = -dot(si.sh_n, dp_du);
    si.sh_s = normalize(V3 { fmadd(si.sh_n.x, d, dp_du.x), fmadd(si.sh_n.y, d, dp_du.y), fmadd(si.sh_n.z, d, dp_du.z) });
    si.sh_t = cross(si.sh_n, si.sh_s);
    si.wi = si.to_local(-ray.d);
    si.shape = tr.shape;
    return si;
}

// interaction.h:156-161
static V3 offset_p(V3 p, V3 n, V3 d) {
    float mag = (1.f + std::fmax(std::fabs(p.x), std::fmax(std::fabs(p.y), std::fabs(p.z)))) * RayEpsilon;
    if (std::signbit(dot(n, d))) mag = -mag;
    return { fmadd(mag, n.x, p.x), fmadd(mag, n.y, p.y), fmadd(mag, n.z, p.z) };
}
static Ray spawn_ray(V3 p, V3 n, V3 d) { Ray r; r.o = offset_p(p, n, d); r.d = d; r.maxt = Largest; return r; }
static Ray spawn_ray_to(V3 p, V3 n, V3 target) {
    Ray r; r.o = offset_p(p, n, target - p);
    V3 d = target - r.o; float dist = norm(d);
    r.d = d / dist; r.maxt = dist * (1.f - ShadowEpsilon);
    return r;
}

// fresnel.h:93-117
static float fresnel_conductor(float ci, float eta_r, float eta_i) {
    float ci2 = ci * ci, si2 = 1.f - ci2, si4 = si2 * si2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - si2;
    float a2pb2 = std::sqrt(std::fmax(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r, 0.f));
    float a = std::sqrt(std::fmax(.5f * (a2pb2 + temp_1), 0.f));
    float term_1 = a2pb2 + ci2, term_2 = 2.f * ci * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a2pb2 * ci2 + si4, term_4 = term_2 * si2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

struct BsdfOut { V3 val { 0, 0, 0 }; float pdf = 0; V3 wo { 0, 0, 0 }; float spdf = 0, seta = 1; bool delta = false, null_ = false; V3 weight { 0, 0, 0 }; };

// BSDF::eval_pdf_sample = eval_pdf(wo) + sample(sample1, sample2)  (include/mitsuba/render/bsdf.h)
static BsdfOut eval_pdf_sample(const Bsdf &b, V3 wi, V3 wo, float s1, V2 s2) {
    BsdfOut o;
    bool flip = b.twosided && wi.z < 0.f;      // twosided.cpp: evaluate the front BRDF with mirrored wi / wo
    if (flip) { wi.z = -wi.z; wo.z = -wo.z; }
    if (b.type == 0) {                          // diffuse.cpp:101-175
        float ci = wi.z, co = wo.z;
        if (ci > 0.f && co > 0.f) { o.val = b.refl * InvPi * co; o.pdf = InvPi * co; }
        if (ci > 0.f) {
            o.wo = square_to_cosine_hemisphere(s2);
            o.spdf = InvPi * o.wo.z; o.seta = 1.f;
            if (o.spdf > 0.f) o.weight = b.refl;
        }
    } else if (b.type == 1) {                   // conductor.cpp:248-310 (eval/pdf of a delta lobe are 0)
        float ci = wi.z;
        if (ci > 0.f) {
            o.wo = { -wi.x, -wi.y, wi.z }; o.spdf = 1.f; o.seta = 1.f; o.delta = true;
            o.weight = { b.refl.x * fresnel_conductor(ci, b.eta.x, b.k.x), b.refl.y * fresnel_conductor(ci, b.eta.y, b.k.y),
                         b.refl.z * fresnel_conductor(ci, b.eta.z, b.k.z) };
        }
    } else {                                    // dielectric.cpp sample (both lobes enabled)
        float ci = wi.z, r_i, ct, eit, eti;
        fresnel(ci, b.eta_s, r_i, ct, eit, eti);
        float t_i = 1.f - r_i;
        bool sel_r = s1 <= r_i;
        o.spdf = sel_r ? r_i : t_i; o.delta = true;
        o.wo = sel_r ? V3 { -wi.x, -wi.y, wi.z } : V3 { -eti * wi.x, -eti * wi.y, ct };
        o.seta = sel_r ? 1.f : eit;
        o.weight = sel_r ? b.refl : b.trans * sqr(eti);
    }
    if (flip) o.wo.z = -o.wo.z;
    return o;
}

static float mis_weight(float a, float b) {     // path.cpp:300-305
    a *= a; b *= b; float w = a / (a + b);
    return std::isfinite(w) ? w : 0.f;
}


// ================================================================================ emitters
struct DS { V3 p, n, d; float dist = 0, pdf = 0; int emitter = -1; };

// rectangle.cpp:166-180 + shape.cpp:358-375 + area.cpp:115-166 (uniform radiance branch)
static bool emitter_sample_direction(const Emitter &e, V3 ref_p, V2 sample, DS &ds, V3 &spec) {
    float sx = sample.x * 2.f - 1.f, sy = sample.y * 2.f - 1.f;
    // to_world.transform_affine(Point3f(sx, sy, 0)): fmadd chain over the matrix columns (ex, ey, -, c)
    ds.p = { fmadd(e.ey.x, sy, fmadd(e.ex.x, sx, e.c.x)), fmadd(e.ey.y, sy, fmadd(e.ex.y, sx, e.c.y)),
             fmadd(e.ey.z, sy, fmadd(e.ex.z, sx, e.c.z)) };
    ds.n = e.n; ds.pdf = e.inv_area;
    ds.d = ds.p - ref_p;
    float dist2 = squared_norm(ds.d);
    ds.dist = std::sqrt(dist2);
    ds.d = ds.d / ds.dist;
    float dp = std::fabs(dot(ds.d, ds.n));
    float x = dist2 / dp;
    ds.pdf *= std::isfinite(x) ? x : 0.f;
    bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
    spec = active ? e.radiance / ds.pdf : V3 { 0, 0, 0 };
    return active;
}

// scene.cpp:227-250 (uniform emitter pmf) + :299-353
static void sample_emitter_direction(const Scene &s, const SI &si, V2 sample, DS &ds, V3 &spec) {
    size_t count = s.emitters.size();
    ds = DS(); spec = { 0, 0, 0 };
    if (count == 0) return;
    int index = 0; float emitter_weight = 1.f, pmf = 1.f;
    if (count > 1) {
        float cf = (float) count, scaled = sample.x * cf;
        index = (int) std::min((uint32_t) scaled, (uint32_t) count - 1u);
        emitter_weight = cf; sample.x = scaled - (float) index;
        pmf = 1.f / cf;       // m_emitter_pmf (scene.cpp:69)
    }
    bool active = emitter_sample_direction(s.emitters[index], si.p, sample, ds, spec);
    ds.emitter = index;
    if (count > 1) { ds.pdf *= pmf; spec = spec * emitter_weight; }
    active = active && ds.pdf != 0.f;
    if (active) {
        Hit h = intersect(s, spawn_ray_to(si.p, si.n, ds.p), true);
        if (h.prim >= 0) { spec = { 0, 0, 0 }; ds.pdf = 0.f; }
    }
}

// scene.cpp:355-366 + area.cpp:168-196 + shape.cpp:377-388
static float pdf_emitter_direction(const Scene &s, const DS &ds) {
    const Emitter &e = s.emitters[ds.emitter];
    float dp = dot(ds.d, ds.n);
    if (!(dp < 0.f)) return 0.f;
    float adp = std::fabs(dp);
    float pdf = e.inv_area * (adp != 0.f ? (ds.dist * ds.dist) / adp : 0.f);
    float pmf = s.emitters.size() > 1 ? 1.f / (float) s.emitters.size() : 1.f;
    return pdf * pmf;
}

// ================================================================================ path.cpp:95-287 (JIT draw order)
static V3 path_sample(const Scene &s, PCG32 &rng, Ray ray, int max_depth, int rr_depth, bool &valid_ray) {
    V3 throughput { 1, 1, 1 }, result { 0, 0, 0 };
    float eta = 1.f; int depth = 0;
    valid_ray = false;
    V3 prev_p { 0, 0, 0 }; float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true;
    if (max_depth == 0) return result;
    bool active = true;
    while (active) {
        Hit h = intersect(s, ray, false);
        SI si = surface_interaction(s, ray, h);
        int em = si.valid() ? s.shape_emitter[si.shape] : -1;
        if (em >= 0) {     // direct emission + MIS (:158-174)
            DS ds; ds.p = si.p; ds.n = si.sh_n; ds.emitter = em;
            V3 rel = si.p - prev_p; ds.dist = norm(rel); ds.d = rel / ds.dist;
            float em_pdf = !prev_bsdf_delta ? pdf_emitter_direction(s, ds) : 0.f;
            float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
            V3 le = (si.wi.z > 0.f && prev_bsdf_pdf > 0.f) ? s.emitters[em].radiance : V3 { 0, 0, 0 };
            le = le * mis_bsdf;
            result = { fmadd(throughput.x, le.x, result.x), fmadd(throughput.y, le.y, result.y), fmadd(throughput.z, le.z, result.z) };
        }
        bool active_next = (depth + 1 < max_depth) && si.valid();
        // every active lane of the recorded loop draws its 6 numbers (2 NEE, 1 + 2 BSDF, 1 RR) before it can stop
        V2 em_sample { 0, 0 }; em_sample.x = rng.next_float32(); em_sample.y = rng.next_float32();
        float sample_1 = rng.next_float32();
        V2 sample_2; sample_2.x = rng.next_float32(); sample_2.y = rng.next_float32();
        float rr_sample = rng.next_float32();
        // path.cpp:256 (JIT: no early break) — valid_ray |= active && si.is_valid(), with the loop mask `active`
        if (si.valid()) valid_ray = true;
        if (!active_next) break;
        const Bsdf &bsdf = s.bsdfs[s.shape_bsdf[si.shape]];
        bool active_em = bsdf.type == 0;          // BSDFFlags::Smooth
        DS ds; V3 em_weight { 0, 0, 0 }, wo { 0, 0, 0 };
        if (active_em) {
            sample_emitter_direction(s, si, em_sample, ds, em_weight);
            active_em = ds.pdf != 0.f;
            wo = si.to_local(ds.d);
        }
        BsdfOut b = eval_pdf_sample(bsdf, si.wi, wo, sample_1, sample_2);
        if (active_em) {
            float mis_em = mis_weight(ds.pdf, b.pdf);
            V3 c = b.val * em_weight * mis_em;
            result = { fmadd(throughput.x, c.x, result.x), fmadd(throughput.y, c.y, result.y), fmadd(throughput.z, c.z, result.z) };
        }
        ray = spawn_ray(si.p, si.n, si.to_world(b.wo));
        throughput = throughput * b.weight;
        eta *= b.seta;
        prev_p = si.p; prev_bsdf_pdf = b.spdf; prev_bsdf_delta = b.delta;
        depth += 1;
        float tmax = std::fmax(throughput.x, std::fmax(throughput.y, throughput.z));
        float rr_prob = std::fmin(tmax * sqr(eta), .95f);
        bool rr_active = depth >= rr_depth, rr_continue = rr_sample < rr_prob;
        if (rr_active) throughput = throughput * rcp(rr_prob);
        active = active_next && (!rr_active || rr_continue) && tmax != 0.f;
    }
    // (a camera ray that escapes or a path cut short leaves valid_ray as is: no environment emitter)
    return valid_ray ? result : V3 { 0, 0, 0 };
}

}  // namespace orc

// ================================================================================ render driver + C entry points
namespace orc {

struct Persp { float s2c[16]; float to_world[16]; float near_clip, far_clip; };

// perspective.cpp sample_ray with the sample->camera matrix composed as in sensor.h:262-303 (in double)
static void persp_setup(Persp &P, float x_fov_deg, int w, int h, int cw, int ch, int cx, int cy) {
    double W = w, H = h, aspect = W / H;
    double relx = cw / W, rely = ch / H, offx = cx / W, offy = cy / H;
    double tanh_ = std::tan(0.5 * (double) x_fov_deg * 3.14159265358979323846 / 180.0), n_ = P.near_clip;
    for (int i = 0; i < 16; ++i) P.s2c[i] = 0.f;
    P.s2c[0] = (float) (relx / -0.5 * tanh_ * n_); P.s2c[3] = (float) ((offx / -0.5 + 1.0) * tanh_ * n_);
    P.s2c[5] = (float) (rely / (-0.5 * aspect) * tanh_ * n_); P.s2c[7] = (float) ((offy / (-0.5 * aspect) + 1.0 / aspect) * tanh_ * n_);
    P.s2c[11] = (float) n_; P.s2c[15] = 1.f;
}
static float persp_sample_ray(const Persp &P, V2 pos, Ray &out) {
    const float *m = P.s2c;
    V3 np { m[0] * pos.x + m[1] * pos.y + m[3], m[4] * pos.x + m[5] * pos.y + m[7], m[8] * pos.x + m[9] * pos.y + m[11] };
    V3 d = normalize(np);
    const float *t = P.to_world;
    out.o = { t[3], t[7], t[11] };
    out.d = { t[0] * d.x + t[1] * d.y + t[2] * d.z, t[4] * d.x + t[5] * d.y + t[6] * d.z, t[8] * d.x + t[9] * d.y + t[10] * d.z };
    float inv_z = rcp(d.z), near_t = P.near_clip * inv_z, far_t = P.far_clip * inv_z;
    out.o = out.o + out.d * near_t; out.maxt = far_t - near_t;
    return 1.f;
}
}  // namespace orc

using namespace orc;
extern "C" {
void orc_block_put(int kind, float param, int size_x, int size_y, int off_x, int off_y, int border, int channels,
                   int coalesce, int normalize, int64_t n, const float *pos, const float *values, double *tensor64);

// flat scene description shared with the product's lrt_scene_desc (include/lensrt.h)
void *orc_scene_create(int nv, const float *verts, int nt, const int *idx, const int *tri_shape, int ns, const int *shape_bsdf,
                       const int *shape_emitter, int nb, const float *bsdf16, int ne, const float *emit16) {
    Scene *s = new Scene();
    (void) nv;
    for (int i = 0; i < nt; ++i) {
        Tri t; const int *f = idx + 3 * i;
        t.p0 = { verts[3 * f[0]], verts[3 * f[0] + 1], verts[3 * f[0] + 2] };
        t.p1 = { verts[3 * f[1]], verts[3 * f[1] + 1], verts[3 * f[1] + 2] };
        t.p2 = { verts[3 * f[2]], verts[3 * f[2] + 1], verts[3 * f[2] + 2] };
        t.shape = tri_shape[i]; s->tris.push_back(t);
    }
    s->shape_bsdf.assign(shape_bsdf, shape_bsdf + ns); s->shape_emitter.assign(shape_emitter, shape_emitter + ns);
    for (int i = 0; i < nb; ++i) {
        const float *r = bsdf16 + 16 * i; Bsdf b;
        b.type = (int) r[0]; b.twosided = r[1] != 0.f; b.refl = { r[2], r[3], r[4] }; b.eta = { r[5], r[6], r[7] };
        b.k = { r[8], r[9], r[10] }; b.trans = { r[11], r[12], r[13] }; b.eta_s = r[14];
        s->bsdfs.push_back(b);
    }
    for (int i = 0; i < ne; ++i) {
        const float *r = emit16 + 16 * i; Emitter e;
        e.radiance = { r[0], r[1], r[2] }; e.shape = (int) r[3]; e.c = { r[4], r[5], r[6] }; e.ex = { r[7], r[8], r[9] };
        e.ey = { r[10], r[11], r[12] };
        // rectangle.cpp:115-123: frame from dp_du = 2 ex, dp_dv = 2 ey; area = |dp_du x dp_dv|
        V3 cr = cross(e.ex * 2.f, e.ey * 2.f);
        e.n = normalize(cross(e.ex, e.ey)); e.inv_area = rcp(norm(cr));
        s->emitters.push_back(e);
    }
    s->build();
    return s;
}
void orc_scene_destroy(void *h) { delete (Scene *) h; }

// closest hit / any hit for n rays: od7 = o, d, maxt -> out4 = t, u, v, prim (prim = -1: miss)
void orc_scene_intersect(void *h, int any_hit, int64_t n, const float *od7, float *out4) {
    const Scene &s = *(Scene *) h;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < n; ++i) {
        const float *r = od7 + 7 * i; Ray ray; ray.o = { r[0], r[1], r[2] }; ray.d = { r[3], r[4], r[5] }; ray.maxt = r[6];
        Hit hit = intersect(s, ray, any_hit != 0);
        float *o = out4 + 4 * i; o[0] = hit.t; o[1] = hit.u; o[2] = hit.v; o[3] = (float) hit.prim;
    }
}

// SamplingIntegrator::render (integrator.cpp:247-330) + render_sample (:419-491), one pass.
// sensor: cam != NULL -> realistic lens (oracle Camera), else perspective with (fov, to_world, clips).
// film: full size (w,h), crop (cw,ch,cx,cy), filter kind/param; storage64 [ch, cw, 4] (R,G,B,W), no border.
// Samples idx in [idx_begin, idx_end) of the wavefront W*H*spp are rendered (sharding hook).
void orc_render(void *scene_h, void *cam_h, float fov, const float *to_world16, float near_clip, float far_clip,
                int w, int h, int cw, int ch, int cx, int cy, int filter, float filter_param, uint32_t seed, int spp,
                int max_depth, int rr_depth, int64_t idx_begin, int64_t idx_end, double *storage64, int64_t *n_lens_miss) {
    const Scene &s = *(Scene *) scene_h;
    const Camera *cam = (const Camera *) cam_h;
    Persp P; P.near_clip = near_clip; P.far_clip = far_clip;
    if (!cam) { std::memcpy(P.to_world, to_world16, 64); persp_setup(P, fov, w, h, cw, ch, cx, cy); }
    float scale_x = 1.f / (float) cw, scale_y = 1.f / (float) ch;
    float offs_x = -(float) cx * scale_x, offs_y = -(float) cy * scale_y;
    bool coalesce = spp >= 4;
    int64_t miss = 0;
    size_t npix = (size_t) cw * ch;
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = 1;
#endif
    (void) nthreads;
#pragma omp parallel reduction(+ : miss)
    {
        std::vector<double> local(npix * 4, 0.0);
#pragma omp for schedule(dynamic, 4096)
        for (int64_t idx = idx_begin; idx < idx_end; ++idx) {
            PCG32 rng; sampler_seed_jit(rng, seed, (uint32_t) idx);
            uint32_t pix = (uint32_t) idx / (uint32_t) spp;
            int py = (int) (pix / (uint32_t) cw), px = (int) (pix - (uint32_t) cw * (uint32_t) py);
            float posx = (float) (px + cx), posy = (float) (py + cy);
            float u0 = rng.next_float32(), u1 = rng.next_float32();
            float spx = posx + u0, spy = posy + u1;
            V2 adjusted { fmadd(spx, scale_x, offs_x), fmadd(spy, scale_y, offs_y) };
            Ray ray; float weight;
            if (cam) {
                V2 ap; ap.x = rng.next_float32(); ap.y = rng.next_float32();
                if (cam->shutter_open_time > 0.f) (void) rng.next_float32();   // the time sample (integrator.cpp:442-444)
                weight = cam->sample_ray(adjusted, ap, ray);
            } else weight = persp_sample_ray(P, adjusted, ray);
            V3 L { 0, 0, 0 };
            if (weight > 0.f) { bool valid; L = path_sample(s, rng, ray, max_depth, rr_depth, valid); }
            else miss++;
            float vals[4] = { weight * L.x, weight * L.y, weight * L.z, 1.f };
            bool box = filter == 0;
            float pp[2] = { box ? posx : spx, box ? posy : spy };
            orc_block_put(filter, filter_param, cw, ch, cx, cy, 0, 4, coalesce, 0, 1, pp, vals, local.data());
        }
#pragma omp critical
        for (size_t i = 0; i < npix * 4; ++i) storage64[i] += local[i];
    }
    if (n_lens_miss) *n_lens_miss = miss;
}
}
