// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). Integrator-side restatement:
//   SamplingIntegrator::render / render_block / render_sample   src/render/integrator.cpp:120-367,369-418,419-491
//     * JIT branch  (one PCG32 stream per sample, coalesced splat)        :247-330
//     * scalar branch (one stream per pixel, Morton order inside spiral blocks, 31-tap filter) :163-245,369-418
//   Spiral::next_block                                          src/render/spiral.cpp:26-75
//   PathIntegrator::sample                                      src/integrators/path.cpp:95-287
//   Scene::sample_emitter_direction / pdf_emitter_direction     src/render/scene.cpp:227-366
//   AreaLight                                                   src/emitters/area.cpp:83-211
//   Shape::sample_direction / pdf_direction                     src/render/shape.cpp:358-388
//   Rectangle (analytic)                                        src/shapes/rectangle.cpp:115-185,449-565
//   Mesh Moeller-Trumbore + compute_surface_interaction         include/mitsuba/render/mesh.h:432-452,
//                                                               src/render/mesh.cpp:1376-1500
//   SurfaceInteraction::initialize_sh_frame / finalize          include/mitsuba/render/interaction.h:245-255,464-484
//   Interaction::spawn_ray / spawn_ray_to / offset_p            include/mitsuba/render/interaction.h:133-162
//   SmoothDiffuse / SmoothConductor / SmoothDielectric          src/bsdfs/{diffuse,conductor,dielectric}.cpp
//   TwoSidedBRDF                                                src/bsdfs/twosided.cpp
//   BitmapTexture (bilinear) + dr::Texture::eval_nonaccel       src/textures/bitmap.cpp:696-711,
//                                                               ext/drjit/include/drjit/texture.h:346-412,1143-1172
//   PerspectiveCamera::sample_ray                               src/sensors/perspective.cpp:201-237
// Geometry model: triangle meshes (with optional vertex normals / texture coordinates) in a plain binary BVH, plus —
// when the caller registers them — ANALYTIC rectangles as the reference intersects them (the product tessellates
// rectangles into two triangles carrying the rectangle's uv parameterisation; both models are available here so that
// this file can be pinned against the reference and then used to check the product).
// Pinned against the reference's own scalar_rgb build: tests/golden/ref_integrator.npz (function level) and
// tests/golden/ref_renders.npz (images, scalar seeding convention); see tests/test_ref_parity.py.
#include "oracle.h"
#include "oracle_film.h"
#include <algorithm>
#include <cstring>
#include <cstdio>

namespace orc {

struct Tri { V3 p0, p1, p2; int shape; int flags; V3 n0, n1, n2; V2 uv0, uv1, uv2; };   // flags: 1 normals, 2 uvs
struct Bsdf { int type; bool twosided; V3 refl, eta, k, trans; float eta_s; int tex; };  // 0 diffuse 1 conductor 2 dielectric
struct EmTri { V3 p0, e0, e1; float pmf, cdf; bool has_n; V3 n0, n1, n2; };
// kind 0 rectangle, 1 mesh, 2 constant environment (c = bounding-sphere centre, inv_area slot unused, bs_radius)
struct Emitter { V3 radiance; int shape; V3 c, ex, ey, n; float inv_area; int kind = 0; std::vector<EmTri> tris; float sum = 0; float bs_radius = 1;
                 bool directional = false; };      // directional: src/emitters/directionalarea.cpp (particle tracer only)
struct Rect { int shape; float M[12], Minv[12]; V3 dp_du, dp_dv, n; };                  // rows of the affine 3x4 matrices
// kind 0 bitmap (wrap: 0 repeat 1 mirror 2 clamp); procedural: 1 checkerboard, 2 checkerlines, 3 comb2d, 4 polarlines
struct Texture { int w, h, wrap; std::vector<float> data; float to_uv[6]; int kind = 0; V3 color0 { 0, 0, 0 }, color1 { 0, 0, 0 }; float linewidth = 0, spokes = 0; };
struct Sphere { int shape; V3 center; float radius; bool flip; float M[12], Minv[12];      // rows of to_world / to_object
                bool disk = false; V3 n { 0, 0, 1 }; };   // disk: src/shapes/disk.cpp — unit disk in the xy plane of object space, frame normal n

struct BNode { V3 lo, hi; int left, right, first, count; };

struct Scene {
    std::vector<Tri> tris;
    std::vector<int> shape_bsdf, shape_emitter;
    std::vector<Bsdf> bsdfs;
    std::vector<Emitter> emitters;
    std::vector<Rect> rects;
    std::vector<Texture> textures;
    std::vector<Sphere> spheres;
    int env = -1;                      // index of the environment emitter in `emitters`
    bool hide_emitters = false;
    std::vector<char> shape_is_rect;   // triangles of these shapes are left out of the BVH
    std::vector<BNode> nodes;
    std::vector<int> order;
    void build();
    int build_rec(int first, int count);
};

static V3 vmin(V3 a, V3 b) { return { std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z) }; }
static V3 vmax(V3 a, V3 b) { return { std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z) }; }
static float comp(V3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }
static V3 fmadd3(V3 a, float b, V3 c) { return { fmadd(a.x, b, c.x), fmadd(a.y, b, c.y), fmadd(a.z, b, c.z) }; }
static V3 fmadd3(V3 a, V3 b, V3 c) { return { fmadd(a.x, b.x, c.x), fmadd(a.y, b.y, c.y), fmadd(a.z, b.z, c.z) }; }

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
    order.clear();
    for (size_t i = 0; i < tris.size(); ++i)
        if (shape_is_rect.empty() || !shape_is_rect[tris[i].shape]) order.push_back((int) i);
    nodes.clear(); nodes.reserve(2 * order.size() + 1);
    if (!order.empty()) build_rec(0, (int) order.size());
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

// Transform::transform_affine(Point) / operator*(Vector) (include/mitsuba/core/transform.h:104-111,131-140), rows M[12]
static V3 xf_point(const float *M, V3 p) {
    V3 r { M[3], M[7], M[11] };
    r = fmadd3(V3 { M[0], M[4], M[8] }, p.x, r); r = fmadd3(V3 { M[1], M[5], M[9] }, p.y, r); r = fmadd3(V3 { M[2], M[6], M[10] }, p.z, r);
    return r;
}
static V3 xf_vector(const float *M, V3 v) {
    V3 r = V3 { M[0], M[4], M[8] } * v.x;
    r = fmadd3(V3 { M[1], M[5], M[9] }, v.y, r); r = fmadd3(V3 { M[2], M[6], M[10] }, v.z, r);
    return r;
}

struct Hit { float t = Infinity, u = 0, v = 0; int prim = -1, rect = -1, sphere = -1; };

// sphere.cpp:493-562 (ray_intersect_preliminary_impl, scalar variants: double precision) with math::solve_quadratic
// (include/mitsuba/core/math.h:358-402)
static bool sphere_intersect(const Sphere &sp, const Ray &ray, float &t_out) {
    double radius = sp.radius, cx = sp.center.x, cy = sp.center.y, cz = sp.center.z;
    double maxt = ray.maxt;
    double lx = (double) ray.o.x - cx, ly = (double) ray.o.y - cy, lz = (double) ray.o.z - cz;
    double dx = ray.d.x, dy = ray.d.y, dz = ray.d.z;
    double plane_t = (-lx * dx + -ly * dy + -lz * dz) / std::sqrt(dx * dx + dy * dy + dz * dz);
    bool no_hit = plane_t == 0 && ray.o.x != sp.center.x && ray.o.y != sp.center.y && ray.o.z != sp.center.z;
    float pt = (float) plane_t;                                  // ray(FloatP(plane_t)): evaluated in single precision
    V3 plane_p = fmadd3(ray.d, pt, ray.o);
    double ox = (double) plane_p.x - cx, oy = (double) plane_p.y - cy, oz = (double) plane_p.z - cz;
    no_hit = no_hit && (std::sqrt(ox * ox + oy * oy + oz * oz) > radius);
    double A = dx * dx + dy * dy + dz * dz, B = 2.0 * (ox * dx + oy * dy + oz * dz), C = (ox * ox + oy * oy + oz * oz) - radius * radius;
    bool linear_case = A == 0.0, valid_linear = linear_case && B != 0.0;
    double x0 = -C / B, x1 = x0;
    double discrim = B * B - 4.0 * A * C;                       // dr::fmsub, two roundings in the pinned build
    bool valid_quadratic = !linear_case && discrim >= 0.0;
    if (valid_quadratic) {
        double sq = std::sqrt(discrim), temp = -0.5 * (B + std::copysign(sq, B));
        double x0p = temp / A, x1p = C / temp;
        x0 = std::fmin(x0p, x1p); x1 = std::fmax(x0p, x1p);
    }
    bool found = valid_linear || valid_quadratic;
    double near_t = x0 + plane_t, far_t = x1 + plane_t;
    bool out_bounds = !(near_t <= maxt && far_t >= 0.0), in_bounds = near_t < 0.0 && far_t > maxt;
    if (!(found && !no_hit && !out_bounds && !in_bounds)) return false;
    t_out = near_t < 0.0 ? (float) far_t : (float) near_t;
    return true;
}

// rectangle.cpp:449-470 (ray_intersect_preliminary_impl) / :472-492 (ray_test_impl)
static bool rect_intersect(const Rect &rc, const Ray &ray, float &t, float &lx, float &ly) {
    V3 o = xf_point(rc.Minv, ray.o), d = xf_vector(rc.Minv, ray.d);
    t = -o.z / d.z;
    lx = fmadd(d.x, t, o.x); ly = fmadd(d.y, t, o.y);   // ray(t) = fmadd(d, t, o)
    return t >= 0.f && t <= ray.maxt && std::fabs(lx) <= 1.f && std::fabs(ly) <= 1.f;
}

static Hit intersect(const Scene &s, const Ray &ray, bool any_hit) {
    Hit best;
    Ray r = ray;
    if (!s.nodes.empty()) {
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
    }
    for (size_t k = 0; k < s.rects.size(); ++k) {
        float t, lx, ly;
        if (rect_intersect(s.rects[k], r, t, lx, ly) && t < best.t) {
            best.t = t; best.u = lx; best.v = ly; best.prim = -1; best.rect = (int) k;
            if (any_hit) return best;
            r.maxt = t;
        }
    }
    for (size_t k = 0; k < s.spheres.size(); ++k) {
        float t;
        if (s.spheres[k].disk) {
            // disk.cpp:392-410 (float32): the ray in object space meets z = 0 inside the unit circle
            const Sphere &dk = s.spheres[k];
            V3 o = xf_point(dk.Minv, r.o), d = xf_vector(dk.Minv, r.d);
            t = -o.z / d.z;
            V3 local = fmadd3(d, t, o);
            if (!(t >= 0.f && t <= r.maxt && local.x * local.x + local.y * local.y <= 1.f)) continue;
            if (t < best.t) {
                best.t = t; best.u = local.x; best.v = local.y; best.prim = -1; best.rect = -1; best.sphere = (int) k;
                if (any_hit) return best;
                r.maxt = t;
            }
            continue;
        }
        if (sphere_intersect(s.spheres[k], r, t) && t < best.t) {
            best.t = t; best.u = best.v = 0.f; best.prim = -1; best.rect = -1; best.sphere = (int) k;
            if (any_hit) return best;
            r.maxt = t;
        }
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
    float t = Infinity; V3 p, n, sh_s, sh_t, sh_n, wi; V2 uv { 0, 0 }; int shape = -1;
    bool valid() const { return t != Infinity; }
    V3 to_local(V3 v) const { return { dot(v, sh_s), dot(v, sh_t), dot(v, sh_n) }; }
    V3 to_world(V3 v) const {   // frame.h: fmadd(n, v.z, fmadd(t, v.y, s * v.x))
        return { fmadd(sh_n.x, v.z, fmadd(sh_t.x, v.y, sh_s.x * v.x)), fmadd(sh_n.y, v.z, fmadd(sh_t.y, v.y, sh_s.y * v.x)),
                 fmadd(sh_n.z, v.z, fmadd(sh_t.z, v.y, sh_s.z * v.x)) };
    }
};

// interaction.h:245-255 initialize_sh_frame + :480 wi
static void finalize_si(SI &si, V3 dp_du, const Ray &ray) {
    float d = -dot(si.sh_n, dp_du);
    si.sh_s = normalize(fmadd3(si.sh_n, d, dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { V3 s, t; coordinate_system(si.sh_n, s, t); si.sh_s = s; }
    si.sh_t = cross(si.sh_n, si.sh_s);
    si.wi = si.to_local(-ray.d);
}

static SI surface_interaction(const Scene &s, const Ray &ray, const Hit &h) {
    SI si;
    if (h.rect >= 0) {          // rectangle.cpp:494-565 (non-differentiable branch)
        const Rect &rc = s.rects[h.rect];
        si.t = h.t;
        V3 p = fmadd3(ray.d, h.t, ray.o);
        V3 tr { rc.M[3], rc.M[7], rc.M[11] };
        float dist = dot(tr - p, rc.n);
        si.p = p + rc.n * dist;
        si.n = si.sh_n = rc.n;
        si.uv = { fmadd(h.u, .5f, .5f), fmadd(h.v, .5f, .5f) };
        si.shape = rc.shape;
        finalize_si(si, rc.dp_du, ray);
        return si;
    }
    if (h.sphere >= 0 && s.spheres[h.sphere].disk) {        // disk.cpp:436-520 (non-differentiable branch)
        const Sphere &dk = s.spheres[h.sphere];
        si.t = h.t;
        V3 p = fmadd3(ray.d, h.t, ray.o);
        V3 tr { dk.M[3], dk.M[7], dk.M[11] };
        float dist = dot(tr - p, dk.n);
        si.p = fmadd3(dk.n, dist, p);
        float rr = std::sqrt(sqr(h.u) + sqr(h.v)), inv_r = rcp(rr);
        float v = atan2_drjit(h.v, h.u) * (.5f * InvPi);
        if (v < 0.f) v += 1.f;
        si.uv = { rr, v };
        float cos_phi = rr != 0.f ? h.u * inv_r : 1.f, sin_phi = rr != 0.f ? h.v * inv_r : 0.f;
        V3 dp_du = xf_vector(dk.M, V3 { cos_phi, sin_phi, 0.f });
        si.n = si.sh_n = dk.n;
        si.shape = dk.shape;
        finalize_si(si, dp_du, ray);
        return si;
    }
    if (h.sphere >= 0) {        // sphere.cpp:609-718 (non-differentiable branch)
        const Sphere &sp = s.spheres[h.sphere];
        si.t = h.t;
        si.sh_n = normalize(fmadd3(ray.d, h.t, ray.o) - sp.center);
        si.p = fmadd3(si.sh_n, sp.radius, sp.center);
        V3 local = xf_point(sp.Minv, si.p);
        float rd_2 = sqr(local.x) + sqr(local.y);
        // unit_angle_z (include/mitsuba/core/vector.h): 2 asin(|v - sign(z) e_z| / 2), mirrored for z < 0
        V3 q { local.x, local.y, local.z - std::copysign(1.f, local.z) };
        float temp = 2.f * asin_drjit(.5f * norm(q)), theta = local.z >= 0.f ? temp : Pi - temp;
        float phi = atan2_drjit(local.y, local.x);
        if (phi < 0.f) phi += 2.f * Pi;
        si.uv = { phi * (.5f * InvPi), theta * InvPi };
        V3 dp_du = xf_vector(sp.M, V3 { -local.y, local.x, 0.f }) * (2.f * Pi);
        (void) rd_2;
        if (sp.flip) si.sh_n = -si.sh_n;
        si.n = si.sh_n;
        si.shape = sp.shape;
        finalize_si(si, dp_du, ray);
        return si;
    }
    if (h.prim < 0) return si;
    // mesh.cpp:1376-1500
    const Tri &tr = s.tris[h.prim];
    float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
    si.p = fmadd3(tr.p0, b0, fmadd3(tr.p1, b1, tr.p2 * b2));
    si.t = h.t;
    V3 dp0 = tr.p1 - tr.p0, dp1 = tr.p2 - tr.p0;
    si.n = normalize(cross(dp0, dp1));     // mesh.h face_normal
    si.uv = { b1, b2 };
    V3 dp_du, dp_dv; coordinate_system(si.n, dp_du, dp_dv);
    if (tr.flags & 2) {
        V2 uv0 = tr.uv0, uv1 = tr.uv1, uv2 = tr.uv2;
        si.uv = { fmadd(uv2.x, b2, fmadd(uv1.x, b1, uv0.x * b0)), fmadd(uv2.y, b2, fmadd(uv1.y, b1, uv0.y * b0)) };
        V2 duv0 { uv1.x - uv0.x, uv1.y - uv0.y }, duv1 { uv2.x - uv0.x, uv2.y - uv0.y };
        float det = fmsub(duv0.x, duv1.y, duv0.y * duv1.x), inv_det = rcp(det);
        if (det != 0.f) {
            // fmsub(duv1.y, dp0, duv0.y * dp1) * inv_det
            dp_du = V3 { fmsub(duv1.y, dp0.x, duv0.y * dp1.x), fmsub(duv1.y, dp0.y, duv0.y * dp1.y), fmsub(duv1.y, dp0.z, duv0.y * dp1.z) } * inv_det;
        }
    }
    si.sh_n = si.n;
    if (tr.flags & 1) {
        V3 n = fmadd3(tr.n2, b2, fmadd3(tr.n1, b1, tr.n0 * b0));
        float il = rsqrt(squared_norm(n));
        si.sh_n = n * il;
    }
    si.shape = tr.shape;
    finalize_si(si, dp_du, ray);
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
    float a2pb2 = std::sqrt(std::fmax(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r, 0.f));   // safe_sqrt
    float a = std::sqrt(std::fmax(.5f * (a2pb2 + temp_1), 0.f));
    float term_1 = a2pb2 + ci2, term_2 = 2.f * ci * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a2pb2 * ci2 + si4, term_4 = term_2 * si2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

// dr::Texture<Float,2>::eval_nonaccel, bilinear (texture.h:346-412) after BitmapTexture::interpolate_3 (bitmap.cpp:696-711)
static int tex_wrap(int pos, int shape, int mode) {   // texture.h:1143-1172
    if (mode == 2) return std::min(std::max(pos, 0), shape - 1);
    int shifted = pos < 0 ? pos + 1 : pos;
    int div = shifted / shape;                        // divisor<int32_t>: truncating division
    int mod = pos - div * shape;
    if (mod < 0) mod += shape;
    if (mode == 1 && !(((div & 1) == 0) ^ (pos < 0))) mod = shape - 1 - mod;
    return mod;
}
// checkerboard.cpp:70-89, checkerlines.cpp / comb2d.cpp / polarlines.cpp eval (the fork's line patterns); colours are constants
static V3 procedural_eval(const Texture &t, V2 uv_in) {
    auto xf = [&](V2 p) { return V2 { fmadd(t.to_uv[1], p.y, fmadd(t.to_uv[0], p.x, t.to_uv[2])), fmadd(t.to_uv[4], p.y, fmadd(t.to_uv[3], p.x, t.to_uv[5])) }; };
    bool m0;
    if (t.kind == 1) {
        V2 uv = xf(uv_in);
        bool mx = uv.x - std::floor(uv.x) > .5f, my = uv.y - std::floor(uv.y) > .5f;
        m0 = mx == my;
    } else if (t.kind == 2 || t.kind == 3) {
        V2 uv = xf(uv_in); uv.x *= 2.f; uv.y *= 2.f;
        uv.x -= std::floor(uv.x); uv.y -= std::floor(uv.y);
        bool mx = std::fabs(2.f * uv.x - 1.f) > 1.f - t.linewidth, my = std::fabs(2.f * uv.y - 1.f) > 1.f - t.linewidth;
        m0 = t.kind == 2 ? (mx || my) : !(mx && my);
    } else {
        V2 uv = xf(V2 { 2.f * uv_in.x - 1.f, 2.f * uv_in.y - 1.f });
        float theta = atan2_drjit(uv.x, uv.y), radius = std::sqrt(uv.x * uv.x + uv.y * uv.y);
        theta = theta / (2.f * Pi) * t.spokes;
        m0 = (std::fabs(2.f * (radius - std::floor(radius)) - 1.f) > 1.f - t.linewidth) ||
             (std::fabs(2.f * (theta - std::floor(theta)) - 1.f) > 1.f - t.linewidth * t.spokes / (2.f * Pi * radius));
    }
    return m0 ? t.color0 : t.color1;
}
static V3 texture_eval(const Texture &t, V2 uv_in) {
    if (t.kind != 0) return procedural_eval(t, uv_in);
    V2 uv { fmadd(t.to_uv[1], uv_in.y, fmadd(t.to_uv[0], uv_in.x, t.to_uv[2])), fmadd(t.to_uv[4], uv_in.y, fmadd(t.to_uv[3], uv_in.x, t.to_uv[5])) };
    float px = fmadd(uv.x, (float) t.w, -.5f), py = fmadd(uv.y, (float) t.h, -.5f);
    int ix = (int) std::floor(px), iy = (int) std::floor(py);
    float w1x = px - (float) ix, w1y = py - (float) iy, w0x = 1.f - w1x, w0y = 1.f - w1y;
    int x0 = tex_wrap(ix, t.w, t.wrap), x1 = tex_wrap(ix + 1, t.w, t.wrap), y0 = tex_wrap(iy, t.h, t.wrap), y1 = tex_wrap(iy + 1, t.h, t.wrap);
    auto px3 = [&](int x, int y) { const float *q = t.data.data() + 3 * ((size_t) y * t.w + x); return V3 { q[0], q[1], q[2] }; };
    V3 out { 0, 0, 0 };
    out = fmadd3(px3(x0, y0), w0x * w0y, out); out = fmadd3(px3(x1, y0), w1x * w0y, out);
    out = fmadd3(px3(x0, y1), w0x * w1y, out); out = fmadd3(px3(x1, y1), w1x * w1y, out);
    return out;
}

struct BsdfOut { V3 val { 0, 0, 0 }; float pdf = 0; V3 wo { 0, 0, 0 }; float spdf = 0, seta = 1; bool delta = false, null_ = false; V3 weight { 0, 0, 0 }; };

static V3 bsdf_reflectance(const Scene &s, const Bsdf &b, V2 uv) {
    return b.tex >= 0 ? texture_eval(s.textures[b.tex], uv) : b.refl;
}

// BSDF::eval_pdf_sample = eval_pdf(wo) + sample(sample1, sample2)  (include/mitsuba/render/bsdf.h)
static BsdfOut eval_pdf_sample(const Scene &s, const Bsdf &b, V2 uv, V3 wi, V3 wo, float s1, V2 s2) {
    BsdfOut o;
    bool flip = b.twosided && wi.z < 0.f;      // twosided.cpp: evaluate the front BRDF with mirrored wi / wo
    if (flip) { wi.z = -wi.z; wo.z = -wo.z; }
    if (b.type == 0) {                          // diffuse.cpp:101-175
        V3 refl = bsdf_reflectance(s, b, uv);
        float ci = wi.z, co = wo.z;
        if (ci > 0.f && co > 0.f) { o.val = refl * InvPi * co; o.pdf = InvPi * co; }   // warp::square_to_cosine_hemisphere_pdf
        if (ci > 0.f) {
            o.wo = square_to_cosine_hemisphere(s2);
            o.spdf = InvPi * o.wo.z; o.seta = 1.f;
            if (o.spdf > 0.f) o.weight = refl;
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
        o.wo = sel_r ? V3 { -wi.x, -wi.y, wi.z } : V3 { -eti * wi.x, -eti * wi.y, ct };   // reflect(wi) / refract(wi, cos_theta_t, eta_ti)
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
// Mesh::sample_position (src/render/mesh.cpp:745-803) with DiscreteDistribution::sample_reuse (core/distr_1d.h:116-182) and
// warp::square_to_uniform_triangle (core/warp.h:153-156)
static void mesh_sample_position(const Emitter &e, V2 sample, V3 &p, V3 &n) {
    float target = sample.y * e.sum;
    size_t lo = 0, hi = e.tris.size() - 1;              // dr::binary_search over [valid.x, valid.y]: first index with !(cdf < target)
    while (lo < hi) { size_t mid = (lo + hi) >> 1; if (e.tris[mid].cdf < target) lo = mid + 1; else hi = mid; }
    const EmTri &t = e.tris[lo];
    float norm = e.inv_area;
    float pmf_n = t.pmf * norm, cdf_prev = lo > 0 ? e.tris[lo - 1].cdf * norm : 0.f;
    float sy = (sample.y - cdf_prev) / pmf_n;
    float tt = std::sqrt(std::fmax(1.f - sample.x, 0.f)), b0 = 1.f - tt, b1 = tt * sy;
    p = fmadd3(t.e0, b0, fmadd3(t.e1, b1, t.p0));
    if (t.has_n) n = fmadd3(t.n0, 1.f - b0 - b1, fmadd3(t.n1, b0, t.n2 * b1));
    else n = cross(t.e0, t.e1);
    n = normalize(n);
}

static bool emitter_sample_direction(const Emitter &e, V3 ref_p, V2 sample, DS &ds, V3 &spec) {
    if (e.kind == 2) {          // constant.cpp:113-142: uniform direction, the point 2 max(R, |p - c|) away
        float z = fnmadd(2.f, sample.y, 1.f), r = std::sqrt(std::fmax(fnmadd(z, z, 1.f), 0.f)), sn, cs;   // warp.h:250-255 (circ)
        sincos(2.f * Pi * sample.x, sn, cs);
        V3 d { r * cs, r * sn, z };
        float radius = std::fmax(e.bs_radius, norm(ref_p - e.c)), dist = 2.f * radius;
        ds.p = fmadd3(d, dist, ref_p); ds.n = -d; ds.d = d; ds.dist = dist;
        ds.pdf = .25f * InvPi;
        spec = e.radiance / ds.pdf;
        return true;
    }
    if (e.kind == 1) mesh_sample_position(e, sample, ds.p, ds.n);
    else if (e.kind == 3) {     // Disk::sample_position (disk.cpp:173-190) under the default Shape::sample_direction
        V2 q = square_to_uniform_disk_concentric(sample);
        ds.p = { fmadd(e.ey.x, q.y, fmadd(e.ex.x, q.x, e.c.x)), fmadd(e.ey.y, q.y, fmadd(e.ex.y, q.x, e.c.y)),
                 fmadd(e.ey.z, q.y, fmadd(e.ex.z, q.x, e.c.z)) };
        ds.n = e.n;
    } else {
        float sx = sample.x * 2.f - 1.f, sy = sample.y * 2.f - 1.f;
        // to_world.transform_affine(Point3f(sx, sy, 0)): fmadd chain over the matrix columns (ex, ey, -, c)
        ds.p = { fmadd(e.ey.x, sy, fmadd(e.ex.x, sx, e.c.x)), fmadd(e.ey.y, sy, fmadd(e.ex.y, sx, e.c.y)),
                 fmadd(e.ey.z, sy, fmadd(e.ex.z, sx, e.c.z)) };
        ds.n = e.n;
    }
    ds.pdf = e.inv_area;
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
static void sample_emitter_direction(const Scene &s, const SI &si, V2 sample, bool test_visibility, DS &ds, V3 &spec) {
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
    if (active && test_visibility) {
        Hit h = intersect(s, spawn_ray_to(si.p, si.n, ds.p), true);
        if (h.prim >= 0 || h.rect >= 0 || h.sphere >= 0) { spec = { 0, 0, 0 }; ds.pdf = 0.f; }
    }
}

// scene.cpp:355-366 + area.cpp:168-196 + shape.cpp:377-388
static float pdf_emitter_direction(const Scene &s, const DS &ds) {
    const Emitter &e = s.emitters[ds.emitter];
    float pmf0 = s.emitters.size() > 1 ? 1.f / (float) s.emitters.size() : 1.f;
    if (e.kind == 2) return (.25f * InvPi) * pmf0;      // constant.cpp:144-149
    float dp = dot(ds.d, ds.n);
    if (!(dp < 0.f)) return 0.f;
    float adp = std::fabs(dp);
    float pdf = e.inv_area * (adp != 0.f ? (ds.dist * ds.dist) / adp : 0.f);
    float pmf = s.emitters.size() > 1 ? 1.f / (float) s.emitters.size() : 1.f;
    return pdf * pmf;
}

// ================================================================================ path.cpp:95-287
// scalar = false: the JIT draw order (every active lane draws its 6 numbers per iteration, no early exit);
// scalar = true : the scalar variant's control flow — `break` before any draw once the path cannot continue
//                 (path.cpp:178-179) and emitter samples drawn only at smooth surfaces (any_or<true> evaluates the mask).
// g_rr_scene (test hook, orc_set_rr_scene): Russian roulette DETACHED as in prb (path.cpp:270-275 dr::detach(rr_prob)) — the
// survival probabilities and the termination test follow the throughput of this frozen copy of the scene, so that finite
// differences of a perturbed scene's render see the same paths with the same roulette weights.
static const Scene *g_rr_scene = nullptr;
static bool g_trace = false;     // orc_set_trace: print every vertex of path_sample (debugging aid for pinning)

static V3 path_sample(const Scene &s, PCG32 &rng, Ray ray, int max_depth, int rr_depth, bool scalar, bool &valid_ray) {
    V3 throughput { 1, 1, 1 }, result { 0, 0, 0 }, throughput_rr { 1, 1, 1 };
    float eta = 1.f; int depth = 0;
    valid_ray = false;
    V3 prev_p { 0, 0, 0 }; float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true;
    if (max_depth == 0) return result;
    if (max_depth < 0) max_depth = 0x7fffffff;         // -1 = unbounded (integrator.cpp:765-790)
    valid_ray = !s.hide_emitters && s.env >= 0;        // path.cpp:113-114
    bool active = true;
    while (active) {
        Hit h = intersect(s, ray, false);
        SI si = surface_interaction(s, ray, h);
        int em = si.valid() ? s.shape_emitter[si.shape] : s.env;     // si.emitter(scene): the environment for a ray that escaped
        if (g_trace) std::printf("depth %d hit %d t %.9g p %.9g %.9g %.9g shape %d n %.6g %.6g %.6g wi %.6g %.6g %.6g thr %.6g %.6g %.6g res %.6g %.6g %.6g\n", depth, (int) si.valid(), si.t,
                                 si.p.x, si.p.y, si.p.z, si.shape, si.n.x, si.n.y, si.n.z, si.wi.x, si.wi.y, si.wi.z, throughput.x, throughput.y, throughput.z, result.x, result.y, result.z);
        if (em >= 0) {     // direct emission + MIS (:158-174)
            DS ds; ds.p = si.p; ds.n = si.sh_n; ds.emitter = em;
            V3 rel = si.p - prev_p; ds.dist = norm(rel); ds.d = rel / ds.dist;
            if (!si.valid()) ds.d = ray.d;                          // records.h: d = -si.wi for an invalid interaction
            float em_pdf = !prev_bsdf_delta ? pdf_emitter_direction(s, ds) : 0.f;
            float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
            V3 le = ((si.wi.z > 0.f || !si.valid()) && prev_bsdf_pdf > 0.f) ? s.emitters[em].radiance : V3 { 0, 0, 0 };
            le = le * mis_bsdf;
            result = fmadd3(throughput, le, result);
        }
        bool active_next = (depth + 1 < max_depth) && si.valid();
        if (scalar && !active_next) break;
        const Bsdf *bsdf = si.valid() ? &s.bsdfs[s.shape_bsdf[si.shape]] : nullptr;
        bool active_em = active_next && bsdf->type == 0;          // BSDFFlags::Smooth
        V2 em_sample { 0, 0 };
        if (!scalar || active_em) { em_sample.x = rng.next_float32(); em_sample.y = rng.next_float32(); }
        float sample_1 = rng.next_float32();
        V2 sample_2; sample_2.x = rng.next_float32(); sample_2.y = rng.next_float32();
        float rr_sample = 0.f;
        if (!scalar) rr_sample = rng.next_float32();
        // path.cpp:256 (JIT: no early break) — valid_ray |= active && si.is_valid(), with the loop mask `active`
        if (si.valid()) valid_ray = true;
        if (!active_next) break;
        DS ds; V3 em_weight { 0, 0, 0 }, wo { 0, 0, 0 };
        if (active_em) {
            sample_emitter_direction(s, si, em_sample, true, ds, em_weight);
            active_em = ds.pdf != 0.f;
            wo = si.to_local(ds.d);
        }
        BsdfOut b = eval_pdf_sample(s, *bsdf, si.uv, si.wi, wo, sample_1, sample_2);
        if (g_trace) std::printf("   em_active %d ds.pdf %.9g emitter %d em_w %.6g %.6g %.6g | bsdf val %.6g %.6g %.6g pdf %.6g wo %.6g %.6g %.6g spdf %.6g w %.6g %.6g %.6g\n", (int) active_em, ds.pdf,
                                 ds.emitter, em_weight.x, em_weight.y, em_weight.z, b.val.x, b.val.y, b.val.z, b.pdf, b.wo.x, b.wo.y, b.wo.z, b.spdf, b.weight.x, b.weight.y, b.weight.z);
        if (active_em) {
            float mis_em = mis_weight(ds.pdf, b.pdf);
            V3 c = b.val * em_weight * mis_em;
            result = fmadd3(throughput, c, result);
        }
        ray = spawn_ray(si.p, si.n, si.to_world(b.wo));
        throughput = throughput * b.weight;
        throughput_rr = g_rr_scene ? throughput_rr * eval_pdf_sample(*g_rr_scene, g_rr_scene->bsdfs[g_rr_scene->shape_bsdf[si.shape]], si.uv, si.wi, wo,
                                                                     sample_1, sample_2).weight : throughput;
        eta *= b.seta;
        prev_p = si.p; prev_bsdf_pdf = b.spdf; prev_bsdf_delta = b.delta;
        depth += 1;
        float tmax = std::fmax(throughput_rr.x, std::fmax(throughput_rr.y, throughput_rr.z));
        float rr_prob = std::fmin(tmax * sqr(eta), .95f);
        if (scalar) rr_sample = rng.next_float32();
        bool rr_active = depth >= rr_depth, rr_continue = rr_sample < rr_prob;
        if (rr_active) { throughput = throughput * rcp(rr_prob); if (g_rr_scene) throughput_rr = throughput_rr * rcp(rr_prob); }
        active = active_next && (!rr_active || rr_continue) && tmax != 0.f;
    }
    return valid_ray ? result : V3 { 0, 0, 0 };
}

}  // namespace orc

// ================================================================================ render driver + C entry points
namespace orc {

struct Persp { float s2c[16]; float to_world[16]; float near_clip, far_clip; };

// perspective.cpp sample_ray with the sample->camera matrix composed as in sensor.h:228-265 (in double)
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

// one render_sample (integrator.cpp:419-491): draws, sensor ray, path, values; returns the splat position
struct RenderCtx {
    const Scene *s; const Camera *cam; Persp P;
    float scale_x, scale_y, offs_x, offs_y; int max_depth, rr_depth; bool box;
};
static bool render_sample(const RenderCtx &C, PCG32 &rng, float posx, float posy, bool scalar, float pp[2], float vals[4]) {
    float u0 = rng.next_float32(), u1 = rng.next_float32();
    float spx = posx + u0, spy = posy + u1;
    V2 adjusted { fmadd(spx, C.scale_x, C.offs_x), fmadd(spy, C.scale_y, C.offs_y) };
    Ray ray; float weight;
    if (C.cam) {
        V2 ap; ap.x = rng.next_float32(); ap.y = rng.next_float32();
        if (C.cam->shutter_open_time > 0.f) (void) rng.next_float32();   // the time sample (integrator.cpp:442-444)
        weight = C.cam->sample_ray(adjusted, ap, ray);
    } else weight = persp_sample_ray(C.P, adjusted, ray);
    V3 L { 0, 0, 0 };
    bool lens_hit = weight > 0.f;
    // the scalar variant calls sample() for every ray (a dead lens ray carries weight 0; it still walks the scene and
    // consumes random numbers); the JIT branch is restated with the same masking the product uses (dead rays are skipped)
    if (lens_hit || scalar) { bool valid; L = path_sample(*C.s, rng, ray, C.max_depth, C.rr_depth, scalar, valid); }
    vals[0] = weight * L.x; vals[1] = weight * L.y; vals[2] = weight * L.z; vals[3] = 1.f;
    pp[0] = C.box ? posx : spx; pp[1] = C.box ? posy : spy;
    return lens_hit;
}
}  // namespace orc

using namespace orc;
extern "C" {

// flat scene description shared with the product's lrt_scene_desc (include/lensrt.h)
//   normals [nv,3] / uvs [nv,2] / shape_flags [ns] (bit 0: normals, bit 1: uvs) may be NULL;
//   textures: ntex x (w, h, wrap) in tex_dims, pixel data RGB float concatenated in tex_data, 6 floats of to_uv each;
//   bsdf16[15] = texture id + 1 of a diffuse reflectance (0: constant).
void *orc_scene_create2(int nv, const float *verts, const float *normals, const float *uvs, int nt, const int *idx, const int *tri_shape,
                        int ns, const int *shape_bsdf, const int *shape_emitter, const int *shape_flags, int nb, const float *bsdf16,
                        int ne, const float *emit16, int ntex, const int *tex_dims, const float *tex_data, const float *tex_to_uv) {
    Scene *s = new Scene();
    (void) nv;
    auto v3 = [&](const float *a, int i) { return V3 { a[3 * i], a[3 * i + 1], a[3 * i + 2] }; };
    for (int i = 0; i < nt; ++i) {
        Tri t; const int *f = idx + 3 * i;
        t.p0 = v3(verts, f[0]); t.p1 = v3(verts, f[1]); t.p2 = v3(verts, f[2]);
        t.shape = tri_shape[i];
        t.flags = shape_flags ? shape_flags[t.shape] : 0;
        if (!normals) t.flags &= ~1;
        if (!uvs) t.flags &= ~2;
        t.n0 = t.n1 = t.n2 = { 0, 0, 0 }; t.uv0 = t.uv1 = t.uv2 = { 0, 0 };
        if (t.flags & 1) { t.n0 = v3(normals, f[0]); t.n1 = v3(normals, f[1]); t.n2 = v3(normals, f[2]); }
        if (t.flags & 2) { t.uv0 = { uvs[2 * f[0]], uvs[2 * f[0] + 1] }; t.uv1 = { uvs[2 * f[1]], uvs[2 * f[1] + 1] }; t.uv2 = { uvs[2 * f[2]], uvs[2 * f[2] + 1] }; }
        s->tris.push_back(t);
    }
    s->shape_bsdf.assign(shape_bsdf, shape_bsdf + ns); s->shape_emitter.assign(shape_emitter, shape_emitter + ns);
    for (int i = 0; i < nb; ++i) {
        const float *r = bsdf16 + 16 * i; Bsdf b;
        b.type = (int) r[0]; b.twosided = r[1] != 0.f; b.refl = { r[2], r[3], r[4] }; b.eta = { r[5], r[6], r[7] };
        b.k = { r[8], r[9], r[10] }; b.trans = { r[11], r[12], r[13] }; b.eta_s = r[14]; b.tex = (int) r[15] - 1;
        s->bsdfs.push_back(b);
    }
    for (int i = 0; i < ne; ++i) {
        const float *r = emit16 + 16 * i; Emitter e;
        e.radiance = { r[0], r[1], r[2] }; e.shape = (int) r[3]; e.c = { r[4], r[5], r[6] }; e.ex = { r[7], r[8], r[9] };
        e.ey = { r[10], r[11], r[12] };
        // rectangle.cpp:115-123: frame from dp_du = 2 ex, dp_dv = 2 ey; area = |dp_du x dp_dv|
        V3 cr = cross(e.ex * 2.f, e.ey * 2.f);
        e.n = normalize(cross(e.ex, e.ey)); e.inv_area = rcp(norm(cr));
        if (r[13] == 2.f) { e.kind = 2; e.shape = -1; e.bs_radius = r[7]; s->env = i; }     // constant environment: c = bounding-sphere centre
        if (r[13] == 1.f) {     // mesh emitter: Mesh::build_pmf (mesh.cpp:430-470), cdf accumulated in double (distr_1d.h:233-262)
            e.kind = 1; double sum = 0.0;
            for (int t = 0; t < nt; ++t) {
                if (tri_shape[t] != e.shape) continue;
                const Tri &tr = s->tris[t];
                EmTri q; q.p0 = tr.p0; q.e0 = tr.p1 - tr.p0; q.e1 = tr.p2 - tr.p0;
                q.pmf = .5f * norm(cross(q.e0, q.e1)); sum += (double) q.pmf; q.cdf = (float) sum;
                q.has_n = (tr.flags & 1) != 0; q.n0 = tr.n0; q.n1 = tr.n1; q.n2 = tr.n2;
                e.tris.push_back(q);
            }
            e.sum = e.tris.back().cdf; e.inv_area = rcp(e.sum);
        }
        if (r[13] == 3.f) {      // area / directionalarea on a `disk`: c, ex, ey = translation and first two columns of to_world; r[15] = sign of the frame normal
            e.kind = 3;
            V3 nn = normalize(cross(e.ex, e.ey)); e.n = r[15] < 0.f ? -nn : nn;
            // Disk::surface_area (disk.cpp:163-166): pi * |dp_du| * h, h = the height of dp_dv over dp_du
            float du = norm(e.ex), dv = norm(e.ey); V3 fs = e.ex * rcp(du), ft = e.ey * rcp(dv);
            float hh = std::sqrt(sqr(dv) - sqr(dot(ft * dv, fs)));
            e.inv_area = rcp(Pi * du * hh);
        }
        e.directional = r[14] == 1.f;
        s->emitters.push_back(e);
    }
    size_t off = 0;
    for (int i = 0; i < ntex; ++i) {
        Texture t; t.w = tex_dims[3 * i]; t.h = tex_dims[3 * i + 1]; t.wrap = tex_dims[3 * i + 2];
        size_t n = (size_t) t.w * t.h * 3;
        t.data.assign(tex_data + off, tex_data + off + n); off += n;
        for (int k = 0; k < 6; ++k) t.to_uv[k] = tex_to_uv[6 * i + k];
        s->textures.push_back(t);
    }
    s->build();
    return s;
}
void *orc_scene_create(int nv, const float *verts, int nt, const int *idx, const int *tri_shape, int ns, const int *shape_bsdf,
                       const int *shape_emitter, int nb, const float *bsdf16, int ne, const float *emit16) {
    return orc_scene_create2(nv, verts, nullptr, nullptr, nt, idx, tri_shape, ns, shape_bsdf, shape_emitter, nullptr, nb, bsdf16, ne, emit16,
                             0, nullptr, nullptr, nullptr);
}
// Intersect these shapes as the reference's analytic rectangles (rectangle.cpp) instead of their two triangles:
// to_world12 / to_object12 = rows of the affine matrices; the normal is the third column of the inverse transpose.
void orc_scene_set_rectangles(void *h, int n, const int *shape_ids, const float *to_world12, const float *to_object12) {
    Scene *s = (Scene *) h;
    s->rects.clear(); s->shape_is_rect.assign(s->shape_bsdf.size(), 0);
    for (int i = 0; i < n; ++i) {
        Rect r; r.shape = shape_ids[i];
        std::memcpy(r.M, to_world12 + 12 * i, 48); std::memcpy(r.Minv, to_object12 + 12 * i, 48);
        r.dp_du = xf_vector(r.M, V3 { 2.f, 0.f, 0.f }); r.dp_dv = xf_vector(r.M, V3 { 0.f, 2.f, 0.f });
        // Transform * Normal: inverse_transpose columns = rows of the inverse (transform.h:146-155)
        V3 n0 = V3 { r.Minv[0], r.Minv[1], r.Minv[2] } * 0.f;
        n0 = fmadd3(V3 { r.Minv[4], r.Minv[5], r.Minv[6] }, 0.f, n0); n0 = fmadd3(V3 { r.Minv[8], r.Minv[9], r.Minv[10] }, 1.f, n0);
        r.n = normalize(n0);
        s->rects.push_back(r); s->shape_is_rect[r.shape] = 1;
    }
    s->build();
}
// Analytic spheres (sphere.cpp): per sphere 8 floats (centre xyz, radius, flip_normals, -, -, -) and the rows of to_world / to_object
void orc_scene_set_spheres(void *h, int n, const int *shape_ids, const float *sph8, const float *to_world12, const float *to_object12) {
    Scene *s = (Scene *) h;
    s->spheres.clear();
    for (int i = 0; i < n; ++i) {
        Sphere sp; sp.shape = shape_ids[i]; const float *q = sph8 + 8 * i;
        sp.center = { q[0], q[1], q[2] }; sp.radius = q[3]; sp.flip = q[4] != 0.f;
        if (q[5] == 1.f) {      // disk: q[0..2] = to_world * Normal3f(0, 0, 1), normalised here as disk.cpp:122 does (float32)
            sp.disk = true; sp.n = normalize(sp.center);
            for (Emitter &e : s->emitters) if (e.kind == 3 && e.shape == sp.shape) e.n = sp.n;
        }
        std::memcpy(sp.M, to_world12 + 12 * i, 48); std::memcpy(sp.Minv, to_object12 + 12 * i, 48);
        s->spheres.push_back(sp);
    }
}
// Procedural textures: kinds[i] != 0 turns texture i into checkerboard (1) / checkerlines (2) / comb2d (3) / polarlines (4);
// params8 = color0 rgb, color1 rgb, linewidth, spokes
void orc_scene_set_procedural(void *h, int n, const int *kinds, const float *params8) {
    Scene *s = (Scene *) h;
    for (int i = 0; i < n && i < (int) s->textures.size(); ++i) {
        Texture &t = s->textures[i]; const float *q = params8 + 8 * i;
        t.kind = kinds[i]; t.color0 = { q[0], q[1], q[2] }; t.color1 = { q[3], q[4], q[5] }; t.linewidth = q[6]; t.spokes = q[7];
    }
}
void orc_set_trace(int on) { g_trace = on != 0; }
void orc_scene_set_hide_emitters(void *h, int hide) { ((Scene *) h)->hide_emitters = hide != 0; }
void orc_scene_destroy(void *h) { delete (Scene *) h; }
// Russian roulette follows this (unperturbed) scene in every later render; NULL restores the attached behaviour
void orc_set_rr_scene(void *h) { g_rr_scene = (const Scene *) h; }

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

// Scene::ray_intersect with the full SurfaceInteraction: out25 = t, p3, n3, sh_n3, sh_s3, sh_t3, uv2, wi3, shape,
// diffuse reflectance at the hit (rgb; 0 for other BSDFs) — the layout of tests/golden/ref_integrator.npz `isect`
void orc_scene_interaction(void *h, int64_t n, const float *od7, float *out25) {
    const Scene &s = *(Scene *) h;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = od7 + 7 * i; Ray ray; ray.o = { r[0], r[1], r[2] }; ray.d = { r[3], r[4], r[5] }; ray.maxt = r[6];
        SI si = surface_interaction(s, ray, intersect(s, ray, false));
        float *o = out25 + 25 * i;
        for (int k = 0; k < 25; ++k) o[k] = 0.f;
        o[0] = si.t; o[21] = -1.f;
        if (!si.valid()) continue;
        const V3 *vs[6] = { &si.p, &si.n, &si.sh_n, &si.sh_s, &si.sh_t };
        for (int k = 0; k < 5; ++k) { o[1 + 3 * k] = vs[k]->x; o[2 + 3 * k] = vs[k]->y; o[3 + 3 * k] = vs[k]->z; }
        o[16] = si.uv.x; o[17] = si.uv.y; o[18] = si.wi.x; o[19] = si.wi.y; o[20] = si.wi.z; o[21] = (float) si.shape;
        const Bsdf &b = s.bsdfs[s.shape_bsdf[si.shape]];
        if (b.type == 0) { V3 c = bsdf_reflectance(s, b, si.uv); o[22] = c.x; o[23] = c.y; o[24] = c.z; }
    }
}

// Scene::sample_emitter_direction(si, sample, test_visibility = false) + pdf_emitter_direction for n reference points:
// out16 = ds.p3, ds.n3, ds.d3, dist, pdf, weight3, pdf_emitter_direction, emitter index
void orc_scene_sample_emitter(void *h, int64_t n, const float *ref_p3, const float *sample2, float *out16) {
    const Scene &s = *(Scene *) h;
    for (int64_t i = 0; i < n; ++i) {
        SI si; si.p = { ref_p3[3 * i], ref_p3[3 * i + 1], ref_p3[3 * i + 2] }; si.n = { 0, 0, 1 };
        DS ds; V3 w;
        sample_emitter_direction(s, si, V2 { sample2[2 * i], sample2[2 * i + 1] }, false, ds, w);
        float *o = out16 + 16 * i;
        o[0] = ds.p.x; o[1] = ds.p.y; o[2] = ds.p.z; o[3] = ds.n.x; o[4] = ds.n.y; o[5] = ds.n.z; o[6] = ds.d.x; o[7] = ds.d.y; o[8] = ds.d.z;
        o[9] = ds.dist; o[10] = ds.pdf; o[11] = w.x; o[12] = w.y; o[13] = w.z;
        o[14] = ds.pdf > 0.f ? pdf_emitter_direction(s, ds) : 0.f; o[15] = ds.pdf > 0.f ? (float) ds.emitter : -1.f;
    }
}

// PathIntegrator::sample (scalar control flow) for n given rays: ray i uses a PCG32 stream seeded like a scalar sampler after
// sampler->seed(seeds[i]) (sampler.cpp:131). out5 = radiance rgb, valid flag, number of random numbers consumed
void orc_path_sample(void *h, int64_t n, const float *od7, const uint32_t *seeds, int max_depth, int rr_depth, float *out5) {
    const Scene &s = *(Scene *) h;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = od7 + 7 * i; Ray ray; ray.o = { r[0], r[1], r[2] }; ray.d = { r[3], r[4], r[5] }; ray.maxt = r[6];
        PCG32 rng; rng.seed((uint64_t) seeds[i], 0xda3e39cb94b95bdbULL);
        PCG32 probe = rng;
        bool valid; V3 L = path_sample(s, rng, ray, max_depth, rr_depth, true, valid);
        int used = 0;
        while (!(probe.state == rng.state) && used < 4096) { (void) probe.next_float32(); ++used; }
        float *o = out5 + 5 * i; o[0] = L.x; o[1] = L.y; o[2] = L.z; o[3] = valid ? 1.f : 0.f; o[4] = (float) used;
    }
}

// BSDF::eval_pdf + BSDF::sample in the local frame for n queries against one 16-float BSDF row:
// out13 = eval rgb, pdf, sampled wo3, sampled pdf, eta, weight rgb, delta flag
void orc_bsdf_eval_pdf_sample(const float *bsdf16, int64_t n, const float *wi3, const float *wo3, const float *s1, const float *s2, float *out13) {
    Scene s; const float *r = bsdf16; Bsdf b;
    b.type = (int) r[0]; b.twosided = r[1] != 0.f; b.refl = { r[2], r[3], r[4] }; b.eta = { r[5], r[6], r[7] };
    b.k = { r[8], r[9], r[10] }; b.trans = { r[11], r[12], r[13] }; b.eta_s = r[14]; b.tex = -1;
    for (int64_t i = 0; i < n; ++i) {
        BsdfOut o = eval_pdf_sample(s, b, V2 { 0, 0 }, V3 { wi3[3 * i], wi3[3 * i + 1], wi3[3 * i + 2] }, V3 { wo3[3 * i], wo3[3 * i + 1], wo3[3 * i + 2] },
                                    s1[i], V2 { s2[2 * i], s2[2 * i + 1] });
        float *q = out13 + 13 * i;
        q[0] = o.val.x; q[1] = o.val.y; q[2] = o.val.z; q[3] = o.pdf; q[4] = o.wo.x; q[5] = o.wo.y; q[6] = o.wo.z; q[7] = o.spdf; q[8] = o.seta;
        q[9] = o.weight.x; q[10] = o.weight.y; q[11] = o.weight.z; q[12] = o.delta ? 1.f : 0.f;
    }
}

// BitmapTexture::eval for n uv pairs (bilinear; wrap 0 repeat / 1 mirror / 2 clamp)
void orc_texture_eval(int w, int h, int wrap, const float *rgb, const float *to_uv6, int64_t n, const float *uv2, float *out3) {
    Texture t; t.w = w; t.h = h; t.wrap = wrap; t.data.assign(rgb, rgb + (size_t) w * h * 3);
    for (int k = 0; k < 6; ++k) t.to_uv[k] = to_uv6 ? to_uv6[k] : (k == 0 || k == 4 ? 1.f : 0.f);
    for (int64_t i = 0; i < n; ++i) {
        V3 c = texture_eval(t, V2 { uv2[2 * i], uv2[2 * i + 1] });
        out3[3 * i] = c.x; out3[3 * i + 1] = c.y; out3[3 * i + 2] = c.z;
    }
}

// PerspectiveCamera::sample_ray for n film samples: out7 = o3, d3, maxt
void orc_persp_sample_ray(float fov, const float *to_world16, float near_clip, float far_clip, int w, int h, int cw, int ch, int cx, int cy,
                          int64_t n, const float *pos2, float *out7) {
    Persp P; P.near_clip = near_clip; P.far_clip = far_clip;
    std::memcpy(P.to_world, to_world16, 64); persp_setup(P, fov, w, h, cw, ch, cx, cy);
    for (int64_t i = 0; i < n; ++i) {
        Ray r; persp_sample_ray(P, V2 { pos2[2 * i], pos2[2 * i + 1] }, r);
        float *o = out7 + 7 * i; o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z; o[6] = r.maxt;
    }
}

static void make_ctx(RenderCtx &C, void *scene_h, void *cam_h, float fov, const float *to_world16, float near_clip, float far_clip,
                     int w, int h, int cw, int ch, int cx, int cy, int filter, int max_depth, int rr_depth) {
    C.s = (const Scene *) scene_h; C.cam = (const Camera *) cam_h;
    C.P.near_clip = near_clip; C.P.far_clip = far_clip;
    if (!C.cam) { std::memcpy(C.P.to_world, to_world16, 64); persp_setup(C.P, fov, w, h, cw, ch, cx, cy); }
    C.scale_x = 1.f / (float) cw; C.scale_y = 1.f / (float) ch;
    C.offs_x = -(float) cx * C.scale_x; C.offs_y = -(float) cy * C.scale_y;
    C.max_depth = max_depth; C.rr_depth = rr_depth; C.box = filter == 0;
}

// SamplingIntegrator::render, JIT branch (integrator.cpp:247-330) + render_sample (:419-491), one pass.
// sensor: cam != NULL -> realistic lens (oracle Camera), else perspective with (fov, to_world, clips).
// film: full size (w,h), crop (cw,ch,cx,cy), filter kind/param; storage64 [ch, cw, 4] (R,G,B,W), no border.
// Samples idx in [idx_begin, idx_end) of the wavefront W*H*spp are rendered (sharding hook).
void orc_render(void *scene_h, void *cam_h, float fov, const float *to_world16, float near_clip, float far_clip,
                int w, int h, int cw, int ch, int cx, int cy, int filter, float filter_param, uint32_t seed, int spp,
                int max_depth, int rr_depth, int64_t idx_begin, int64_t idx_end, double *storage64, int64_t *n_lens_miss) {
    RenderCtx C; make_ctx(C, scene_h, cam_h, fov, to_world16, near_clip, far_clip, w, h, cw, ch, cx, cy, filter, max_depth, rr_depth);
    Filter flt; flt.init(filter, filter_param);
    bool coalesce = spp >= 4;
    int64_t miss = 0;
    size_t npix = (size_t) cw * ch;
#pragma omp parallel reduction(+ : miss)
    {
        std::vector<double> local(npix * 4, 0.0);
#pragma omp for schedule(dynamic, 4096)
        for (int64_t idx = idx_begin; idx < idx_end; ++idx) {
            PCG32 rng; sampler_seed_jit(rng, seed, (uint32_t) idx);
            uint32_t pix = (uint32_t) idx / (uint32_t) spp;
            int py = (int) (pix / (uint32_t) cw), px = (int) (pix - (uint32_t) cw * (uint32_t) py);
            float pp[2], vals[4];
            if (!render_sample(C, rng, (float) (px + cx), (float) (py + cy), false, pp, vals)) miss++;
            block_put(flt, cw, ch, cx, cy, 0, 4, coalesce, false, false, pp[0], pp[1], vals, local.data());
        }
#pragma omp critical
        for (size_t i = 0; i < npix * 4; ++i) storage64[i] += local[i];
    }
    if (n_lens_miss) *n_lens_miss = miss;
}

// SamplingIntegrator::render, SCALAR branch (integrator.cpp:163-245) + render_block (:369-418): blocks of block_size^2
// pixels handed out along the spiral (spiral.cpp:26-75), pixels of a block in Morton order, ONE PCG32 stream per pixel
// seeded with  seed * W * H + block_id * block_size^2 + i  (initseq = PCG32_DEFAULT_STREAM), its spp samples consecutive;
// non-coalesced splat with the 31-tap discretised filter (rfilter.h:66-76). The per-block ImageBlock + put_block merge is
// a pure sum, so samples are splatted straight into the (borderless) film storage.
static uint32_t compact_bits(uint32_t x) {   // gather the even bits
    x &= 0x55555555u; x = (x | (x >> 1)) & 0x33333333u; x = (x | (x >> 2)) & 0x0f0f0f0fu;
    x = (x | (x >> 4)) & 0x00ff00ffu; x = (x | (x >> 8)) & 0x0000ffffu;
    return x;
}
void orc_render_scalar(void *scene_h, void *cam_h, float fov, const float *to_world16, float near_clip, float far_clip,
                       int w, int h, int cw, int ch, int cx, int cy, int filter, float filter_param, uint32_t base_seed, uint32_t seed,
                       int spp, int max_depth, int rr_depth, int block_size, double *storage64, int64_t *n_lens_miss) {
    RenderCtx C; make_ctx(C, scene_h, cam_h, fov, to_world16, near_clip, far_clip, w, h, cw, ch, cx, cy, filter, max_depth, rr_depth);
    Filter flt; flt.init(filter, filter_param);
    // Spiral (spiral.cpp): block ids in spiral order starting at the centre block
    int bx = (cw + block_size - 1) / block_size, by = (ch + block_size - 1) / block_size, n_blocks = bx * by;
    std::vector<int> block_x(n_blocks), block_y(n_blocks);
    {
        int px = bx / 2, py = by / 2, dir = 0, steps_left = 1, spiral_size = 1;   // Right, Down, Left, Up
        for (int counter = 0; counter < n_blocks; ++counter) {
            block_x[counter] = px; block_y[counter] = py;
            if (counter + 1 == n_blocks) break;
            do {
                switch (dir) { case 0: ++px; break; case 1: ++py; break; case 2: --px; break; default: --py; break; }
                if (--steps_left == 0) {
                    dir = (dir + 1) % 4;
                    if (dir == 2 || dir == 0) ++spiral_size;
                    steps_left = spiral_size;
                }
            } while (px < 0 || py < 0 || px >= bx || py >= by);
        }
    }
    uint32_t seed_eff = seed * (uint32_t) (cw * ch);            // integrator.cpp:196 (film_size = crop size)
    int64_t miss = 0;
    size_t npix = (size_t) cw * ch;
#pragma omp parallel reduction(+ : miss)
    {
        std::vector<double> local(npix * 4, 0.0);
#pragma omp for schedule(dynamic, 1)
        for (int b = 0; b < n_blocks; ++b) {
            uint32_t block_id = (uint32_t) b;                   // single pass: m_passes_left - 1 = 0
            uint32_t bseed = seed_eff + block_id * (uint32_t) (block_size * block_size);
            int ox = block_x[b] * block_size, oy = block_y[b] * block_size;
            int sx = std::min(block_size, cw - ox), sy = std::min(block_size, ch - oy);
            for (uint32_t i = 0; i < (uint32_t) (block_size * block_size); ++i) {
                PCG32 rng; rng.seed((uint64_t) (base_seed + bseed + i), 0xda3e39cb94b95bdbULL);   // sampler.cpp:131
                uint32_t mx = compact_bits(i), my = compact_bits(i >> 1);
                if ((int) mx >= sx || (int) my >= sy) continue;
                float posx = (float) ((int) mx + ox + cx), posy = (float) ((int) my + oy + cy);
                for (int j = 0; j < spp; ++j) {
                    float pp[2], vals[4];
                    if (!render_sample(C, rng, posx, posy, true, pp, vals)) miss++;
                    block_put(flt, cw, ch, cx, cy, 0, 4, false, false, true, pp[0], pp[1], vals, local.data());
                }
            }
        }
#pragma omp critical
        for (size_t i = 0; i < npix * 4; ++i) storage64[i] += local[i];
    }
    if (n_lens_miss) *n_lens_miss = miss;
}
}

#include "oracle_ptracer_impl.h"   // the particle tracer (uses the static helpers above)
