// lensrt — surface interaction, BSDFs, area emitters and the per-bounce logic of the path integrator.
// Replaces: Mesh::compute_surface_interaction (src/render/mesh.cpp:1376-1500, flat-shaded triangles),
// SurfaceInteraction::initialize_sh_frame / finalize (include/mitsuba/render/interaction.h:245-255,464-484),
// Interaction::spawn_ray / spawn_ray_to / offset_p (:133-162), coordinate_system (core/vector.h:116-136),
// SmoothDiffuse (src/bsdfs/diffuse.cpp:101-175), SmoothConductor (conductor.cpp:248-310 + fresnel_conductor
// render/fresnel.h:93-117), SmoothDielectric (dielectric.cpp sample), TwoSidedBRDF (twosided.cpp),
// AreaLight (src/emitters/area.cpp:83-211) on rectangles (src/shapes/rectangle.cpp:166-185),
// Shape::sample_direction / pdf_direction (src/render/shape.cpp:358-388),
// Scene::sample_emitter_direction / pdf_emitter_direction (src/render/scene.cpp:227-366),
// vertex normals / texture coordinates of Mesh::compute_surface_interaction (src/render/mesh.cpp:1448-1500),
// BitmapTexture bilinear lookups (src/textures/bitmap.cpp:696-711 -> dr::Texture::eval_nonaccel,
// ext/drjit/include/drjit/texture.h:346-412, wrap modes :1143-1172).
#pragma once
#include "lrt_trace.cuh"
#include "lrt_lens.cuh"

namespace lrt {

enum : int { BSDF_DIFFUSE = 0, BSDF_CONDUCTOR = 1, BSDF_DIELECTRIC = 2 };

// kind 0: RGB float texels, row-major; wrap 0 repeat 1 mirror 2 clamp. kinds 1..4: checkerboard, checkerlines, comb2d, polarlines
struct TexView { const float *data; int w, h, wrap, kind; float to_uv[6]; float color0[3], color1[3], linewidth, spokes; };

struct SceneView {
    BvhView bvh;
    const int *shape_bsdf, *shape_emitter;
    const float4 *bsdfs;      // 4 x float4: [type, twosided, refl.rgb | eta.rgb, k.rgb | trans.rgb, eta_s, texture id + 1]
    const float4 *emitters;   // 5 x float4: [radiance.rgb, shape | c.xyz, inv_area | ex.xyz | ey.xyz | n.xyz]
    int n_emitters;
    // optional per-triangle attributes, indexed like bvh.tris (4 x float4: n0.xyz n1.x | n1.yz n2.xy | n2.z uv0.xy uv1.x |
    // uv1.y uv2.xy -); the triangle's flag word (third 128-bit word, .w) says which are valid: bit 0 normals, bit 1 uvs
    // (bits 2 / 3 mark the two triangles of a tessellated rectangle, see make_surface)
    const float4 *tri_attr;
    const TexView *textures;
    // area emitters on general meshes (src/render/mesh.cpp:430-470,745-803): per emitter triangle 6 x float4 =
    // [p0.xyz, cdf | e0.xyz, pmf | e1.xyz, has_normals | n0 | n1 | n2], in the mesh's own face order; the emitter record
    // (second float4) then holds (first triangle, count, area sum, 1 / area sum) and its last float4 .w = 1 marks the kind
    const float4 *em_tris;
    // analytic spheres (sphere.cpp), outside the BVH: 8 x float4 = [c.xyz, r | shape, flip, -, - | to_world rows | to_object rows].
    // A hit on sphere k is reported as primitive n_tris + k.
    const float4 *spheres;
    int n_spheres, n_tris;
    int env;                  // row of the `constant` environment emitter, -1 = none
};

struct Surface {
    float3 p, n;         // position, geometric normal
    float3 sn, s, t;     // shading frame (sn = interpolated vertex normal when the mesh has them, else n)
    float3 wi;           // incident direction in the local frame
    float2 uv;
    int shape;
    LRT_D float3 to_local(float3 v) const { return make_float3(dot(v, s), dot(v, t), dot(v, sn)); }
    LRT_D float3 to_world(float3 v) const { return s * v.x + t * v.y + sn * v.z; }
};

LRT_D void coordinate_system(float3 n, float3 &s, float3 &t) {
    float sign = copysignf(1.f, n.z);
    float a = -1.f / (sign + n.z), b = n.x * n.y * a;
    s = make_float3(copysignf(1.f, n.z) * (n.x * n.x * a) + 1.f, copysignf(1.f, n.z) * b, n.z >= 0.f ? -n.x : n.x);
    t = make_float3(b, n.y * (n.y * a) + sign, -n.y);
}

// Cold path of make_surface, kept out of line so that the render megakernel's hot loop stays inside the instruction
// cache (scenes without vertex attributes never call it): mesh.cpp:1448-1500 — texture coordinates give si.uv and dp_du,
// vertex normals the shading normal.
static __device__ __noinline__ void surface_attributes(const float4 *__restrict__ ap, int flags, float b0, float b1, float b2, float3 dp0, float3 dp1,
                                                float3 &dp_du, float3 &sn, float2 &uv) {
    const float4 a2 = __ldg(ap + 2);
    if (flags & 2) {
        const float4 a3 = __ldg(ap + 3);
        const float2 uv0 = make_float2(a2.y, a2.z), uv1 = make_float2(a2.w, a3.x), uv2 = make_float2(a3.y, a3.z);
        uv = make_float2(uv2.x * b2 + (uv1.x * b1 + uv0.x * b0), uv2.y * b2 + (uv1.y * b1 + uv0.y * b0));
        const float2 duv0 = make_float2(uv1.x - uv0.x, uv1.y - uv0.y), duv1 = make_float2(uv2.x - uv0.x, uv2.y - uv0.y);
        const float det = duv0.x * duv1.y - duv0.y * duv1.x;
        if (det != 0.f) dp_du = (dp0 * duv1.y - dp1 * duv0.y) * (1.f / det);
    }
    if (flags & 1) {
        const float4 a0 = __ldg(ap), a1 = __ldg(ap + 1);
        const float3 n0 = make_float3(a0.x, a0.y, a0.z), n1 = make_float3(a0.w, a1.x, a1.y), n2 = make_float3(a1.z, a1.w, a2.x);
        sn = normalize(n2 * b2 + (n1 * b1 + n0 * b0));
    }
}

// ---- Dr.Jit's single-precision atan2 / asin (ext/drjit/include/drjit/math.h:519-594, 366-412): minimax / CEPHES polynomials
// in Estrin form. The spherical parameterisation and the `polarlines` pattern branch on their results.
LRT_D float atan2_drjit(float y, float x) {
    const float abs_x = fabsf(x), abs_y = fabsf(y), min_val = fminf(abs_y, abs_x), max_val = fmaxf(abs_x, abs_y);
    const float scaled_min = min_val / max_val, z = scaled_min * scaled_min, z2 = z * z;
    const float a0 = z * -0.33326497518773606976f + 0.99999934166683966009f, a1 = z * -0.13486708938456973185f + 0.19881342388439013552f,
                a2 = z * -0.037006525670417265220f + 0.083863120428809689910f, a3 = 0.0078613793713198150252f;
    const float b0 = z2 * a1 + a0, b1 = z2 * a3 + a2;
    float t = ((z2 * z2) * b1 + b0) * scaled_min;
    if (abs_y > abs_x) t = kPi * .5f - t;
    if (x < 0.f) t = kPi - t;
    const float r = y < 0.f ? -t : t;
    return max_val != 0.f ? r : 0.f;
}
LRT_D float asin_drjit(float x) {
    const float xa = fabsf(x), x2 = x * x;
    const bool big = xa > .5f;
    const float x1 = .5f * (1.f - xa), x3 = big ? x1 : x2, x4 = big ? sqrtf(x1) : xa, y2 = x3 * x3;
    const float a0 = x3 * 7.4953002686e-2f + 1.6666752422e-1f, a1 = x3 * 2.4181311049e-2f + 4.5470025998e-2f, a2 = 4.2163199048e-2f;
    float z1 = (y2 * y2) * a2 + (y2 * a1 + a0);
    z1 = z1 * (x3 * x4) + x4;
    const float r = big ? (.5f * kPi) - (z1 + z1) : z1;
    return copysignf(r, x);
}

// Sphere::ray_intersect_preliminary / ray_test (src/shapes/sphere.cpp:493-607) over the scene's analytic spheres, double
// precision with every operation rounded separately as in the reference's scalar build (math::solve_quadratic,
// include/mitsuba/core/math.h:358-402). Returns the index of the closest sphere hit in [0, maxt] (any_hit: the first found),
// -1 if none. Out of line: scenes without spheres never come here.
static __device__ __noinline__ int spheres_intersect(const SceneView &sc, float3 o, float3 d, float maxt_f, bool any_hit, float &t_out) {
    int best = -1;
    for (int k = 0; k < sc.n_spheres; ++k) {
        const float4 s0 = __ldg(sc.spheres + 8 * (size_t) k);
        if (__ldg(sc.spheres + 8 * (size_t) k + 1).z == 1.f) {
            // `disk` (disk.cpp:392-410, float32): the ray in object space meets z = 0 inside the unit circle
            const float4 *sp = sc.spheres + 8 * (size_t) k;
            const float4 m0 = __ldg(sp + 5), m1 = __ldg(sp + 6), m2 = __ldg(sp + 7);
            auto xf = [](float4 m, float3 v, float w) {        // transform_affine: column by column, two roundings per step
                float r = w != 0.f ? __fadd_rn(__fmul_rn(m.x, v.x), m.w) : __fmul_rn(m.x, v.x);
                r = __fadd_rn(__fmul_rn(m.y, v.y), r);
                return __fadd_rn(__fmul_rn(m.z, v.z), r);
            };
            const float oz = xf(m2, o, 1.f), dz = xf(m2, d, 0.f);
            const float t = __fdiv_rn(-oz, dz);
            const float lx = __fadd_rn(__fmul_rn(xf(m0, d, 0.f), t), xf(m0, o, 1.f)), ly = __fadd_rn(__fmul_rn(xf(m1, d, 0.f), t), xf(m1, o, 1.f));
            if (!(t >= 0.f && t <= maxt_f && __fadd_rn(__fmul_rn(lx, lx), __fmul_rn(ly, ly)) <= 1.f)) continue;
            if (best < 0 || t < t_out) { best = k; t_out = t; maxt_f = fminf(maxt_f, t); }
            if (any_hit) break;
            continue;
        }
        const double radius = s0.w, cx = s0.x, cy = s0.y, cz = s0.z, maxt = maxt_f;
        const double lx = __dsub_rn((double) o.x, cx), ly = __dsub_rn((double) o.y, cy), lz = __dsub_rn((double) o.z, cz);
        const double dx = d.x, dy = d.y, dz = d.z;
        auto dot3 = [](double ax, double ay, double az, double bx, double by, double bz) {
            return __dadd_rn(__dadd_rn(__dmul_rn(ax, bx), __dmul_rn(ay, by)), __dmul_rn(az, bz));
        };
        const double A = dot3(dx, dy, dz, dx, dy, dz);
        const double plane_t = __ddiv_rn(dot3(-lx, -ly, -lz, dx, dy, dz), __dsqrt_rn(A));
        bool no_hit = plane_t == 0.0 && o.x != s0.x && o.y != s0.y && o.z != s0.z;
        const float pt = (float) plane_t;                      // ray(FloatP(plane_t)) is evaluated in single precision
        const double ox = __dsub_rn((double) strict::fmadd(d.x, pt, o.x), cx), oy = __dsub_rn((double) strict::fmadd(d.y, pt, o.y), cy),
                     oz = __dsub_rn((double) strict::fmadd(d.z, pt, o.z), cz);
        const double oo = dot3(ox, oy, oz, ox, oy, oz);
        no_hit = no_hit && __dsqrt_rn(oo) > radius;
        const double B = __dmul_rn(2.0, dot3(ox, oy, oz, dx, dy, dz)), C = __dsub_rn(oo, __dmul_rn(radius, radius));
        const bool linear_case = A == 0.0, valid_linear = linear_case && B != 0.0;
        double x0 = __ddiv_rn(-C, B), x1 = x0;
        const double discrim = __dsub_rn(__dmul_rn(B, B), __dmul_rn(__dmul_rn(4.0, A), C));
        const bool valid_quadratic = !linear_case && discrim >= 0.0;
        if (valid_quadratic) {
            const double sq = __dsqrt_rn(discrim), temp = __dmul_rn(-0.5, __dadd_rn(B, copysign(sq, B)));
            const double x0p = __ddiv_rn(temp, A), x1p = __ddiv_rn(C, temp);
            x0 = fmin(x0p, x1p); x1 = fmax(x0p, x1p);
        }
        const double near_t = __dadd_rn(x0, plane_t), far_t = __dadd_rn(x1, plane_t);
        const bool out_bounds = !(near_t <= maxt && far_t >= 0.0), in_bounds = near_t < 0.0 && far_t > maxt;
        if (!((valid_linear || valid_quadratic) && !no_hit && !out_bounds && !in_bounds)) continue;
        const float t = near_t < 0.0 ? (float) far_t : (float) near_t;
        if (best < 0 || t < t_out) { best = k; t_out = t; maxt_f = fminf(maxt_f, t); }
        if (any_hit) break;
    }
    return best;
}

static __device__ __noinline__ void sphere_surface(const SceneView &sc, int k, float3 ray_o, float3 ray_d, float t, Surface &si);

// kAttr = false compiles the vertex-attribute path out (scenes without normals / uvs / textures / spheres: the fast kernel
// variant). ray_o is only needed for hits on analytic spheres (primitive index >= n_tris).
template <bool kAttr = true, bool kAnalytic = kAttr> LRT_D Surface make_surface(const SceneView &sc, const RayHit &h, float3 ray_o, float3 ray_d) {
    if (kAnalytic && h.tri >= sc.n_tris) { Surface ss; sphere_surface(sc, h.tri - sc.n_tris, ray_o, ray_d, h.t, ss); return ss; }
    const float4 *tp = sc.bvh.tris + 3 * (size_t) h.tri;
    float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
    float3 p0 = make_float3(a.x, a.y, a.z), p1 = make_float3(a.w, b.x, b.y), p2 = make_float3(b.z, b.w, c.x);
    Surface si;
    float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
    si.p = p0 * b0 + (p1 * b1 + p2 * b2);
    const float3 dp0 = p1 - p0, dp1 = p2 - p0;
    // the face normal's cross product with separately rounded products: a component that cancels exactly (a planar shape whose
    // normal lies in a coordinate plane) must come out as the reference's exact 0 — with a fused multiply-add it would be the
    // product's rounding error, of either sign, and coordinate_system() below branches on the sign of n.z
    si.n = normalize(make_float3(strict::sub(strict::mul(dp0.y, dp1.z), strict::mul(dp0.z, dp1.y)), strict::sub(strict::mul(dp0.z, dp1.x), strict::mul(dp0.x, dp1.z)),
                                 strict::sub(strict::mul(dp0.x, dp1.y), strict::mul(dp0.y, dp1.x))));
    float3 dp_du, dp_dv; coordinate_system(si.n, dp_du, dp_dv);
    si.sn = si.n; si.uv = make_float2(b1, b2);
    const int flags = __float_as_int(c.w);
    // tessellated `rectangle` shapes (bits 2 / 3: first / second triangle of the quad (0,1,2),(0,2,3)): dp_du of the
    // rectangle's own parameterisation (rectangle.cpp:118-121) is an edge of the triangle — the value the uv formula of
    // mesh.cpp:1471-1480 yields for uv = (x, y) / 2 + 1/2, without fetching any attribute
    if (flags & 12) dp_du = (flags & 4) ? dp0 : dp0 - dp1;
    if (kAttr && (flags & 3)) surface_attributes(sc.tri_attr + 4 * (size_t) h.tri, flags, b0, b1, b2, dp0, dp1, dp_du, si.sn, si.uv);
    si.s = normalize(dp_du - si.sn * dot(si.sn, dp_du));
    si.t = cross(si.sn, si.s);
    si.wi = si.to_local(-ray_d);
    si.shape = __float_as_int(c.y);
    return si;
}

// Sphere::compute_surface_interaction (sphere.cpp:609-718, non-differentiable branch) + initialize_sh_frame
static __device__ __noinline__ void sphere_surface(const SceneView &sc, int k, float3 ray_o, float3 ray_d, float t, Surface &si) {
    const float4 *sp = sc.spheres + 8 * (size_t) k;
    const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
    const float3 c = make_float3(s0.x, s0.y, s0.z);
    if (s1.z == 1.f) {        // `disk`: Disk::compute_surface_interaction (disk.cpp:436-520, non-differentiable branch)
        const float4 w0 = __ldg(sp + 2), w1 = __ldg(sp + 3), w2 = __ldg(sp + 4), m0 = __ldg(sp + 5), m1 = __ldg(sp + 6);
        const float3 p = ray_o + ray_d * t, n = c;               // c: the frame normal
        si.p = p + n * dot(make_float3(w0.w, w1.w, w2.w) - p, n);      // re-projected onto the disk's plane
        const float3 od = make_float3(m0.z * ray_d.z + (m0.y * ray_d.y + m0.x * ray_d.x), m1.z * ray_d.z + (m1.y * ray_d.y + m1.x * ray_d.x), 0.f);
        const float lx = od.x * t + (m0.z * ray_o.z + (m0.y * ray_o.y + (m0.x * ray_o.x + m0.w))),
                    ly = od.y * t + (m1.z * ray_o.z + (m1.y * ray_o.y + (m1.x * ray_o.x + m1.w)));
        const float r = sqrtf(lx * lx + ly * ly), inv_r = 1.f / r;
        float v = atan2_drjit(ly, lx) * (.5f * kInvPi);
        if (v < 0.f) v += 1.f;
        si.uv = make_float2(r, v);
        const float cp = r != 0.f ? lx * inv_r : 1.f, sn = r != 0.f ? ly * inv_r : 0.f;
        const float3 dp_du = make_float3(w0.y * sn + w0.x * cp, w1.y * sn + w1.x * cp, w2.y * sn + w2.x * cp);
        si.n = si.sn = n;
        si.s = normalize(dp_du - n * dot(n, dp_du));
        si.t = cross(n, si.s);
        si.wi = si.to_local(-ray_d);
        si.shape = (int) s1.x;
        return;
    }
    float3 n = normalize((ray_o + ray_d * t) - c);
    si.p = c + n * s0.w;                                         // re-projected onto the sphere
    const float4 w0 = __ldg(sp + 2), w1 = __ldg(sp + 3), w2 = __ldg(sp + 4), m0 = __ldg(sp + 5), m1 = __ldg(sp + 6), m2 = __ldg(sp + 7);
    const float3 local = make_float3(m0.z * si.p.z + (m0.y * si.p.y + (m0.x * si.p.x + m0.w)), m1.z * si.p.z + (m1.y * si.p.y + (m1.x * si.p.x + m1.w)),
                                     m2.z * si.p.z + (m2.y * si.p.y + (m2.x * si.p.x + m2.w)));
    const float3 q = make_float3(local.x, local.y, local.z - copysignf(1.f, local.z));       // unit_angle_z (drjit/sphere.h:73-82)
    const float temp = 2.f * asin_drjit(.5f * sqrtf(dot(q, q))), theta = local.z >= 0.f ? temp : kPi - temp;
    float phi = atan2_drjit(local.y, local.x);
    if (phi < 0.f) phi += 2.f * kPi;
    si.uv = make_float2(phi * (.5f * kInvPi), theta * kInvPi);
    const float ux = -local.y, uy = local.x;                     // dp_du = to_world * (-y, x, 0) * 2 pi
    const float3 dp_du = make_float3(w0.y * uy + w0.x * ux, w1.y * uy + w1.x * ux, w2.y * uy + w2.x * ux) * (2.f * kPi);
    if (s1.y != 0.f) n = -n;
    si.n = si.sn = n;
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { float3 tt; coordinate_system(n, si.s, tt); }
    else si.s = normalize(dp_du - n * dot(n, dp_du));
    si.t = cross(n, si.s);
    si.wi = si.to_local(-ray_d);
    si.shape = (int) s1.x;
}

LRT_D float3 offset_p(float3 p, float3 n, float3 d) {
    float mag = (1.f + fmaxf(fabsf(p.x), fmaxf(fabsf(p.y), fabsf(p.z)))) * kRayEpsilon;
    mag = copysignf(mag, dot(n, d));
    return p + n * mag;
}

LRT_D float fresnel_conductor(float ci, float eta_r, float eta_i) {
    float ci2 = ci * ci, si2 = 1.f - ci2, si4 = si2 * si2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - si2;
    float a2pb2 = sqrtf(fmaxf(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r, 0.f));
    float a = sqrtf(fmaxf(.5f * (a2pb2 + temp_1), 0.f));
    float term_1 = a2pb2 + ci2, term_2 = 2.f * ci * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a2pb2 * ci2 + si4, term_4 = term_2 * si2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}

// fresnel.h:35-84 (relaxed arithmetic; the lens uses the strict copy in lrt_lens.cuh)
LRT_D void fresnel_dielectric(float ci, float eta, float &r, float &ct, float &eta_it, float &eta_ti) {
    bool outside = ci >= 0.f;
    float rcp_eta = 1.f / eta;
    eta_it = outside ? eta : rcp_eta; eta_ti = outside ? rcp_eta : eta;
    float ct2 = 1.f - (1.f - ci * ci) * (eta_ti * eta_ti);
    float cia = fabsf(ci), cta = sqrtf(fmaxf(ct2, 0.f));
    float a_s = (cia - eta_it * cta) / (cia + eta_it * cta);
    float a_p = (cta - eta_it * cia) / (cta + eta_it * cia);
    r = 0.5f * (a_s * a_s + a_p * a_p);
    if (eta == 1.f) r = 0.f; else if (cia == 0.f) r = 1.f;
    ct = ci >= 0.f ? -cta : cta;
}

// dr::Texture<Float, 2>::wrap (texture.h:1143-1172): integer texel coordinate -> [0, shape)
LRT_D int tex_wrap(int pos, int shape, int mode) {
    if (mode == 2) return min(max(pos, 0), shape - 1);
    const int shifted = pos < 0 ? pos + 1 : pos, div = shifted / shape;
    int mod = pos - div * shape;
    if (mod < 0) mod += shape;
    if (mode == 1 && !(((div & 1) == 0) != (pos < 0))) mod = shape - 1 - mod;
    return mod;
}

// checkerboard.cpp:70-89 and the fork's checkerlines / comb2d / polarlines patterns (constant colours)
static __device__ __noinline__ float3 procedural_eval(const TexView *tv, float2 uv_in) {
    const float *m = tv->to_uv;
    const int kind = tv->kind;
    bool m0;
    if (kind == 4) {
        const float x = 2.f * uv_in.x - 1.f, y = 2.f * uv_in.y - 1.f;
        const float u = m[1] * y + (m[0] * x + m[2]), v = m[4] * y + (m[3] * x + m[5]);
        const float radius = sqrtf(u * u + v * v), theta = atan2_drjit(u, v) / (2.f * kPi) * tv->spokes;
        m0 = (fabsf(2.f * (radius - floorf(radius)) - 1.f) > 1.f - tv->linewidth) ||
             (fabsf(2.f * (theta - floorf(theta)) - 1.f) > 1.f - tv->linewidth * tv->spokes / (2.f * kPi * radius));
    } else {
        float u = m[1] * uv_in.y + (m[0] * uv_in.x + m[2]), v = m[4] * uv_in.y + (m[3] * uv_in.x + m[5]);
        if (kind == 1) m0 = (u - floorf(u) > .5f) == (v - floorf(v) > .5f);
        else {
            u *= 2.f; v *= 2.f; u -= floorf(u); v -= floorf(v);
            const bool mx = fabsf(2.f * u - 1.f) > 1.f - tv->linewidth, my = fabsf(2.f * v - 1.f) > 1.f - tv->linewidth;
            m0 = kind == 2 ? (mx || my) : !(mx && my);
        }
    }
    const float *c = m0 ? tv->color0 : tv->color1;
    return make_float3(c[0], c[1], c[2]);
}

// BitmapTexture::eval for an RGB texture, bilinear filter (out of line: cold for untextured scenes)
static __device__ __noinline__ float3 texture_eval(const TexView *tv, float2 uv_in) {
    if (tv->kind != 0) return procedural_eval(tv, uv_in);
    const float *m = tv->to_uv;
    const float u = m[1] * uv_in.y + (m[0] * uv_in.x + m[2]), v = m[4] * uv_in.y + (m[3] * uv_in.x + m[5]);
    const int w = tv->w, h = tv->h, mode = tv->wrap;
    const float px = fmaf(u, (float) w, -.5f), py = fmaf(v, (float) h, -.5f);
    const float fx = floorf(px), fy = floorf(py);
    const int ix = (int) fx, iy = (int) fy;
    const float w1x = px - fx, w1y = py - fy, w0x = 1.f - w1x, w0y = 1.f - w1y;
    const int x0 = tex_wrap(ix, w, mode), x1 = tex_wrap(ix + 1, w, mode), y0 = tex_wrap(iy, h, mode), y1 = tex_wrap(iy + 1, h, mode);
    const float *d = tv->data;
    auto texel = [&](int x, int y) { const float *q = d + 3 * ((size_t) y * w + x); return make_float3(__ldg(q), __ldg(q + 1), __ldg(q + 2)); };
    float3 out = texel(x0, y0) * (w0x * w0y);
    out += texel(x1, y0) * (w1x * w0y); out += texel(x0, y1) * (w0x * w1y); out += texel(x1, y1) * (w1x * w1y);
    return out;
}

struct BsdfResult {
    float3 val; float pdf;          // eval_pdf(wo)
    float3 wo; float spdf, seta;    // sample
    float3 weight; bool delta, transmitted;   // transmitted: the dielectric's refraction lobe was sampled
};

// BSDF::eval_pdf_sample: eval/pdf for the emitter direction `wo` and a fresh sample from (s1, s2)
template <bool kAttr = true> LRT_D BsdfResult bsdf_eval_pdf_sample(const SceneView &sc, int bsdf_id, float2 uv, float3 wi, float3 wo, float s1, float s2x, float s2y) {
    const float4 *bp = sc.bsdfs + 4 * (size_t) bsdf_id;
    float4 b0 = __ldg(bp), b1 = __ldg(bp + 1), b2 = __ldg(bp + 2), b3 = __ldg(bp + 3);
    int type = (int) b0.x; bool twosided = b0.y != 0.f;
    float3 refl = make_float3(b0.z, b0.w, b1.x);
    BsdfResult o;
    o.val = make_float3(0, 0, 0); o.pdf = 0.f; o.wo = make_float3(0, 0, 0); o.spdf = 0.f; o.seta = 1.f;
    o.weight = make_float3(0, 0, 0); o.delta = false; o.transmitted = false;
    bool flip = twosided && wi.z < 0.f;
    if (flip) { wi.z = -wi.z; wo.z = -wo.z; }
    if (type == BSDF_DIFFUSE) {
        if (kAttr && b3.w != 0.f) refl = texture_eval(sc.textures + ((int) b3.w - 1), uv);   // `reflectance` is a bitmap texture
        float ci = wi.z, co = wo.z;
        if (ci > 0.f && co > 0.f) { o.val = refl * (kInvPi * co); o.pdf = kInvPi * co; }
        if (ci > 0.f) {
            float px, py; square_to_disk_concentric(s2x, s2y, px, py);   // same (strict) warp as the aperture
            float z = sqrtf(fmaxf(1.f - (px * px + py * py), 0.f));
            o.wo = make_float3(px, py, z); o.spdf = kInvPi * z;
            if (o.spdf > 0.f) o.weight = refl;
        }
    } else if (type == BSDF_CONDUCTOR) {
        float ci = wi.z; o.delta = true;
        if (ci > 0.f) {
            o.wo = make_float3(-wi.x, -wi.y, wi.z); o.spdf = 1.f;
            o.weight = make_float3(refl.x * fresnel_conductor(ci, b1.y, b2.x), refl.y * fresnel_conductor(ci, b1.z, b2.y),
                                   refl.z * fresnel_conductor(ci, b1.w, b2.z));
        }
    } else {
        float ci = wi.z, r_i, ct, eit, eti;
        fresnel_dielectric(ci, b3.z, r_i, ct, eit, eti);
        bool sel_r = s1 <= r_i;
        o.delta = true; o.spdf = sel_r ? r_i : 1.f - r_i;
        o.wo = sel_r ? make_float3(-wi.x, -wi.y, wi.z) : make_float3(-eti * wi.x, -eti * wi.y, ct);
        o.seta = sel_r ? 1.f : eit; o.transmitted = !sel_r;
        o.weight = sel_r ? refl : make_float3(b2.w, b3.x, b3.y) * (eti * eti);
    }
    if (flip) o.wo.z = -o.wo.z;
    return o;
}

LRT_D bool bsdf_is_smooth(const SceneView &sc, int bsdf_id) { return (int) __ldg(sc.bsdfs + 4 * (size_t) bsdf_id).x == BSDF_DIFFUSE; }

struct EmitterSample { float3 p, n, d; float dist, pdf; float3 weight; };

// Mesh::sample_position (mesh.cpp:745-803): face by the area distribution with sample reuse (distr_1d.h:116-182), uniform
// point on it (warp.h:153-156), interpolated vertex normal when the mesh has normals. Out of line: cold for rectangle lights.
static __device__ __noinline__ void mesh_emitter_sample_position(const SceneView &sc, float4 e1, float sx, float sy, float3 &p, float3 &n) {
    const int first = __float_as_int(e1.x), count = __float_as_int(e1.y);
    const float sum = e1.z, norm = e1.w;
    const float target = sy * sum;
    int lo = 0, hi = count - 1;                       // first index whose cdf is not < target (all faces have area > 0)
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(sc.em_tris + 6 * (size_t) (first + mid)).w < target) lo = mid + 1; else hi = mid;
    }
    const float4 *t = sc.em_tris + 6 * (size_t) (first + lo);
    const float4 t0 = __ldg(t), t1 = __ldg(t + 1), t2 = __ldg(t + 2);
    const float cdf_prev = lo > 0 ? __ldg(sc.em_tris + 6 * (size_t) (first + lo - 1)).w * norm : 0.f;
    const float sy2 = (sy - cdf_prev) / (t1.w * norm);
    const float tt = sqrtf(fmaxf(1.f - sx, 0.f)), b0 = 1.f - tt, b1 = tt * sy2;
    const float3 p0 = make_float3(t0.x, t0.y, t0.z), e0 = make_float3(t1.x, t1.y, t1.z), ee1 = make_float3(t2.x, t2.y, t2.z);
    p = e0 * b0 + (ee1 * b1 + p0);
    if (t2.w != 0.f) {
        const float4 a = __ldg(t + 3), b = __ldg(t + 4), c = __ldg(t + 5);
        n = make_float3(a.x, a.y, a.z) * (1.f - b0 - b1) + (make_float3(b.x, b.y, b.z) * b0 + make_float3(c.x, c.y, c.z) * b1);
    } else n = cross(e0, ee1);
    n = normalize(n);
}

// Scene::sample_emitter_direction (scene.cpp:299-353) without the visibility test (done by the caller)
template <bool kAttr = true, bool kAnalytic = kAttr> LRT_D EmitterSample sample_emitter_direction(const SceneView &sc, float3 ref_p, float sx, float sy) {
    EmitterSample es; es.pdf = 0.f; es.weight = make_float3(0, 0, 0); es.p = es.n = es.d = make_float3(0, 0, 0); es.dist = 0.f;
    int count = sc.n_emitters;
    if (count == 0) return es;
    int index = 0; float emitter_weight = 1.f, pmf = 1.f;
    if (count > 1) {
        float cf = (float) count, scaled = sx * cf;
        index = min((int) scaled, count - 1);
        emitter_weight = cf; sx = scaled - (float) index; pmf = 1.f / cf;
    }
    const float4 *ep = sc.emitters + 5 * (size_t) index;
    float4 e0 = __ldg(ep), e1 = __ldg(ep + 1), e2 = __ldg(ep + 2), e3 = __ldg(ep + 3), e4 = __ldg(ep + 4);
    float inv_area = e1.w;
    if (kAnalytic && e4.w == 2.f) {     // `constant` environment (constant.cpp:113-142): uniform direction, the point 2 max(R, |p - c|) away
        const float z = 1.f - 2.f * sy, r = sqrtf(fmaxf(1.f - z * z, 0.f));
        float sn, cs; sincos_drjit(2.f * kPi * sx, sn, cs);
        es.d = make_float3(r * cs, r * sn, z);
        const float3 rel = ref_p - make_float3(e1.x, e1.y, e1.z);
        es.dist = 2.f * fmaxf(e1.w, sqrtf(dot(rel, rel)));
        es.p = ref_p + es.d * es.dist; es.n = -es.d;
        es.weight = make_float3(e0.x, e0.y, e0.z) * (emitter_weight * (4.f * kPi));
        es.pdf = (.25f * kInvPi) * pmf;
        return es;
    }
    if (kAttr && e4.w == 1.f) mesh_emitter_sample_position(sc, e1, sx, sy, es.p, es.n);
    else if (kAnalytic && e4.w == 3.f) {      // on a `disk`: Disk::sample_position (disk.cpp:173-190), concentric warp
        float qx, qy; square_to_disk_concentric(sx, sy, qx, qy);
        es.p = make_float3(e1.x, e1.y, e1.z) + make_float3(e2.x, e2.y, e2.z) * qx + make_float3(e3.x, e3.y, e3.z) * qy;
        es.n = make_float3(e4.x, e4.y, e4.z);
    } else {
        float ux = sx * 2.f - 1.f, uy = sy * 2.f - 1.f;
        es.p = make_float3(e1.x, e1.y, e1.z) + make_float3(e2.x, e2.y, e2.z) * ux + make_float3(e3.x, e3.y, e3.z) * uy;
        es.n = make_float3(e4.x, e4.y, e4.z);
    }
    float3 d = es.p - ref_p;
    float dist2 = dot(d, d);
    es.dist = sqrtf(dist2);
    es.d = d * (1.f / es.dist);
    float dn = dot(es.d, es.n);
    float x = dist2 / fabsf(dn);
    float pdf = inv_area * (isfinite(x) ? x : 0.f);
    bool active = dn < 0.f && pdf != 0.f;
    if (active) {
        es.weight = make_float3(e0.x, e0.y, e0.z) * (emitter_weight / pdf);
        es.pdf = pdf * pmf;
    }
    return es;
}

// Scene::pdf_emitter_direction for a surface hit on emitter `em` seen from `ref_p`
LRT_D float pdf_emitter_direction(const SceneView &sc, int em, float3 d, float dist, float3 n) {
    float dp = dot(d, n);
    if (!(dp < 0.f)) return 0.f;
    float inv_area = __ldg(sc.emitters + 5 * (size_t) em + 1).w;
    float adp = fabsf(dp);
    float pdf = inv_area * (adp != 0.f ? dist * dist / adp : 0.f);
    return sc.n_emitters > 1 ? pdf / (float) sc.n_emitters : pdf;
}

// the environment seen by a ray that escaped (constant.cpp:144-149): uniform over the sphere of directions
LRT_D float pdf_environment(const SceneView &sc) {
    const float pdf = .25f * kInvPi;
    return sc.n_emitters > 1 ? pdf / (float) sc.n_emitters : pdf;
}

// closest hit over the BVH and the analytic spheres (callers outside the render megakernel's resumable traversal)
template <bool kAttr = true> LRT_D RayHit scene_intersect(const SceneView &sc, float3 o, float3 d, float maxt, bool any_hit, unsigned &work) {
    int sph = -1; float sph_t = maxt;
    if (kAttr && sc.n_spheres > 0) {
        sph = spheres_intersect(sc, o, d, maxt, any_hit, sph_t);
        if (sph >= 0 && any_hit) { RayHit h; h.t = sph_t; h.u = h.v = 0.f; h.tri = sc.n_tris + sph; return h; }
        if (sph >= 0) maxt = sph_t;
    }
    RayHit h = bvh_intersect(sc.bvh, o, d, maxt, any_hit, work);
    if (kAttr && h.tri < 0 && sph >= 0) { h.t = sph_t; h.u = h.v = 0.f; h.tri = sc.n_tris + sph; }
    return h;
}

LRT_D float mis_weight(float a, float b) {
    a *= a; b *= b; float w = a / (a + b);
    return isfinite(w) ? w : 0.f;
}

}  // namespace lrt
