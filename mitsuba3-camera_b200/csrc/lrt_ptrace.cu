// lensrt — the particle tracer: light paths from the emitters, connected to the realistic lens at every vertex.
// Replaces ParticleTracerIntegrator (src/integrators/ptracer.cpp:77-371: sample, sample_visible_emitters, prepare_ray,
// trace_light_ray, connect_sensor) under AdjointIntegrator::render (src/render/integrator.cpp:544-760), with
// Scene::sample_emitter / sample_emitter_ray (src/render/scene.cpp:227-296), AreaLight::sample_ray / sample_position
// (src/emitters/area.cpp:92-115,210-245), ConstantBackgroundEmitter::sample_ray (src/emitters/constant.cpp:93-111),
// RealisticLensCamera::sample_direction (src/sensors/realisticlens.cpp:2301-2394) in world space and the normalised,
// non-coalesced ImageBlock::put (src/render/imageblock.cpp:264-379) with weight 0: the image is the R,G,B channels of the
// accumulator as they are (particle-tracer films are not divided by a weight).
//
// One thread walks one light path (rng_mode 0: one PCG32 stream per sample, the JIT variants' convention) and splats every
// connection with global float atomics; rng_mode 1 is the SCALAR AdjointIntegrator's convention — the samples are cut into ranges
// of max(total / (4 n_threads), 1), range k is ONE stream seeded with seed * (total / grain) + k, one thread per range — a parity
// mode whose images can be compared with a scalar_rgb build of the reference (n_threads = that build's thread count).
// Draw order follows the reference as GCC compiles it: bsdf->sample(ctx, si, next_1d(), next_2d()) draws the 2D sample first
// (function arguments are evaluated right to left), and a masked draw of an inactive lane does not advance the generator.
#include "lrt_internal.h"
#include "lrt_rng.cuh"
#include "lrt_shade.cuh"
#include "lrt_scene.h"
#include "lrt_spot_adj.cuh"

namespace lrt {

struct PtArgs {
    SceneView sc;
    const LensSystem *lens;
    FilmParams film;
    float to_world[12], to_camera[12];     // rows of the sensor's to_world and of its inverse
    uint32_t seed; int scalar;
    long long n_total, grain, n_ranges;    // scalar convention: ranges of `grain` samples
    int shard_index, shard_count;
    int max_depth, rr_depth, hide_emitters, shutter_open;
    // perspective sensor (persp != 0; src/sensors/perspective.cpp:283-386): camera -> sample projection, image rectangle at z = 1,
    // its normalisation, crop resolution, clip planes
    int persp; float c2s[16], rect_min[2], rect_max[2], persp_norm, persp_res[2], near_clip, far_clip;
    float sample_scale;
    float *accum;
    unsigned long long *stats;             // [0] light paths, [1] connections splatted, [2] rays traced
    // lens adjoint (k_ptrace<true>): d(loss)/d(accum) in, d(loss)/d(interface table) out
    const float *grad_accum; float *grad_rows;
};

struct PtVertex {
    float3 p, n, s, t, sn, wi; float2 uv; int shape;   // shape -1: not on a surface (environment)
    LRT_D float3 to_local(float3 v) const { return make_float3(dot(v, s), dot(v, t), dot(v, sn)); }
    LRT_D float3 to_world(float3 v) const { return s * v.x + t * v.y + sn * v.z; }
};

struct SensorDS { bool valid; float u, v; float3 p, d; float dist, pdf, weight; float3 ref; float a, b; };   // ref: camera-space point

// Transform4f::transform_affine (transform.h:104-111): translation column first, then one multiply-add per column, strict
LRT_D float3 xf_point_strict(const float *M, float3 p) {
    using strict::fmadd;
    float x = M[3], y = M[7], z = M[11];
    x = fmadd(M[0], p.x, x); y = fmadd(M[4], p.x, y); z = fmadd(M[8], p.x, z);
    x = fmadd(M[1], p.y, x); y = fmadd(M[5], p.y, y); z = fmadd(M[9], p.y, z);
    x = fmadd(M[2], p.z, x); y = fmadd(M[6], p.z, y); z = fmadd(M[10], p.z, z);
    return make_float3(x, y, z);
}

// Sensor::sample_direction, world space in and out (the camera-space part is lens_sample_direction, lrt_lens.cuh)
// PerspectiveCamera::sample_direction + importance (perspective.cpp:283-321,337-386): the pinhole needs no aperture sample
static __device__ __noinline__ SensorDS persp_sample_direction(const PtArgs &A, float3 p_world) {
    using namespace strict;
    SensorDS r; r.valid = false; r.u = r.v = 0.f; r.p = r.d = make_float3(0, 0, 0); r.dist = r.pdf = r.weight = 0.f; r.a = r.b = 0.f;
    const float3 ref = xf_point_strict(A.to_camera, p_world);
    r.ref = ref;
    if (!(ref.z >= A.near_clip && ref.z <= A.far_clip)) return r;
    const float *m = A.c2s;
    const float x = fmadd(m[2], ref.z, fmadd(m[1], ref.y, fmadd(m[0], ref.x, m[3]))), y = fmadd(m[6], ref.z, fmadd(m[5], ref.y, fmadd(m[4], ref.x, m[7]))),
                w = fmadd(m[14], ref.z, fmadd(m[13], ref.y, fmadd(m[12], ref.x, m[15])));
    const float iw = rcp(w), u = mul(x, iw), v = mul(y, iw);
    if (!(u >= 0.f && u <= 1.f && v >= 0.f && v <= 1.f)) return r;
    const float dist = sqrt(add(add(mul(ref.x, ref.x), mul(ref.y, ref.y)), mul(ref.z, ref.z))), inv_dist = rcp(dist);
    const float lx = mul(ref.x, inv_dist), ly = mul(ref.y, inv_dist), lz = mul(ref.z, inv_dist);
    const float inv_ct = rcp(lz), px = mul(lx, inv_ct), py = mul(ly, inv_ct);
    const bool inside = lz > 0.f && px >= A.rect_min[0] && px <= A.rect_max[0] && py >= A.rect_min[1] && py <= A.rect_max[1];
    const float importance = inside ? mul(mul(mul(A.persp_norm, inv_ct), inv_ct), inv_ct) : 0.f;
    r.valid = true; r.u = mul(u, A.persp_res[0]); r.v = mul(v, A.persp_res[1]);
    r.p = xf_point_strict(A.to_world, make_float3(0, 0, 0));
    r.dist = dist;
    r.d = make_float3(mul(sub(r.p.x, p_world.x), inv_dist), mul(sub(r.p.y, p_world.y), inv_dist), mul(sub(r.p.z, p_world.z), inv_dist));
    r.pdf = 1.f; r.weight = mul(mul(importance, inv_dist), inv_dist);
    return r;
}

LRT_D SensorDS sensor_sample_direction(const PtArgs &A, const LensSystem &L, float3 p_world, float a, float b) {
    if (A.persp) return persp_sample_direction(A, p_world);
    SensorDS r; r.valid = false; r.u = r.v = 0.f; r.p = r.d = make_float3(0, 0, 0); r.dist = r.pdf = r.weight = 0.f;
    const float3 ref = xf_point_strict(A.to_camera, p_world);
    r.ref = ref; r.a = a; r.b = b;
    const DirectionSample ds = lens_sample_direction(L, L.ifs, strict::v3(ref.x, ref.y, ref.z), a, b);
    if (!ds.valid) return r;
    r.valid = true; r.u = ds.uv_x; r.v = ds.uv_y;
    r.p = xf_point_strict(A.to_world, make_float3(ds.p.x, ds.p.y, ds.p.z));
    r.dist = ds.dist;
    const float inv = strict::rcp(ds.dist);
    r.d = make_float3(strict::mul(strict::sub(r.p.x, p_world.x), inv), strict::mul(strict::sub(r.p.y, p_world.y), inv),
                      strict::mul(strict::sub(r.p.z, p_world.z), inv));
    r.pdf = ds.pdf; r.weight = ds.weight;
    return r;
}

LRT_D void splat(const PtArgs &A, float u, float v, float3 value, unsigned &n_splat) {
    const FilmParams &F = A.film;
    Footprint fp;
    const float px = u + (float) F.off_x, py = v + (float) F.off_y;     // sensor_ds.uv + block->offset()
    const bool ok = A.scalar ? film_footprint<true>(F, px, py, fp) : film_footprint<false>(F, px, py, fp);
    if (!ok) return;
    const int sw = F.size_x + 2 * F.border, sh = F.size_y + 2 * F.border;
    n_splat++;
    for (int iy = 0; iy < fp.ny; ++iy) {
        const int y = fp.y0 + iy;
        if ((unsigned) y >= (unsigned) sh) continue;
        for (int ix = 0; ix < fp.nx; ++ix) {
            const int x = fp.x0 + ix;
            if ((unsigned) x >= (unsigned) sw) continue;
            const float w = fp.wx[ix] * fp.wy[iy];
            if (w == 0.f) continue;
            float *dst = A.accum + ((size_t) y * sw + x) * 4;
            if (value.x != 0.f) atomicAdd(dst, value.x * w);
            if (value.y != 0.f) atomicAdd(dst + 1, value.y * w);
            if (value.z != 0.f) atomicAdd(dst + 2, value.z * w);
        }
    }
}

// BSDF::eval(ctx, si, wo) for the connection: only the diffuse lobe is non-zero (delta lobes evaluate to 0)
LRT_D float3 bsdf_eval(const SceneView &sc, int bsdf_id, float2 uv, float3 wi, float3 wo) {
    const float4 *bp = sc.bsdfs + 4 * (size_t) bsdf_id;
    const float4 b0 = __ldg(bp), b1 = __ldg(bp + 1), b3 = __ldg(bp + 3);
    if (b0.y != 0.f && wi.z < 0.f) { wi.z = -wi.z; wo.z = -wo.z; }
    if ((int) b0.x != BSDF_DIFFUSE || !(wi.z > 0.f && wo.z > 0.f)) return make_float3(0, 0, 0);
    float3 refl = make_float3(b0.z, b0.w, b1.x);
    if (b3.w != 0.f) refl = texture_eval(sc.textures + ((int) b3.w - 1), uv);
    return refl * (kInvPi * wo.z);
}

// ptracer.cpp:285-371. bsdf_id < 0: a directly visible emitter (no scattering). `weight` comes WITHOUT the sensor's importance
// (sds.weight), which the forward multiplies in and the lens adjoint differentiates.
// connect_sensor (ptracer.cpp:286-371) in pieces around its shadow ray. connect_prepare: is a connection worth a shadow ray, and
// the ray (Interaction::spawn_ray_to, interaction.h:139-146). connect_value: what the connection carries once it is known to be
// unoccluded (nothing in it depends on the visibility).
LRT_D bool connect_prepare(const PtVertex &v, const SensorDS &sds, float3 weight, float3 &so, float3 &sd, float &smax) {
    if (!(sds.valid && sds.pdf > 0.f && sds.weight != 0.f && (weight.x != 0.f || weight.y != 0.f || weight.z != 0.f))) return false;
    so = offset_p(v.p, v.n, sds.p - v.p);
    sd = sds.p - so; const float sdist = sqrtf(dot(sd, sd)); sd = sd * (1.f / sdist);
    smax = sdist * (1.f - kShadowEpsilon);
    return true;
}
LRT_D float3 connect_value(const PtArgs &A, const PtVertex &v, int bsdf_id, float3 weight, float3 sd) {
    const bool has_bsdf = bsdf_id >= 0;
    float3 surface_weight = make_float3(1, 1, 1);
    const float3 local_d = v.to_local(sd);
    if (v.shape >= 0) {
        if (!has_bsdf) surface_weight *= fmaxf(0.f, local_d.z);
        else {
            const float wi_dot_geo_n = dot(v.n, v.to_world(v.wi)), wo_dot_geo_n = dot(v.n, sd);
            const bool valid = (wi_dot_geo_n * v.wi.z > 0.f) && (wo_dot_geo_n * local_d.z > 0.f);
            const float correction = valid ? fabsf((v.wi.z * wo_dot_geo_n) / (local_d.z * wi_dot_geo_n)) : 0.f;
            surface_weight = bsdf_eval(A.sc, bsdf_id, v.uv, v.wi, local_d) * correction;
        }
    } else if (!has_bsdf && local_d.z <= 0.f) surface_weight = make_float3(0, 0, 0);
    return weight * surface_weight;
}
template <bool kAdj>
LRT_D void connect_sensor(const PtArgs &A, const LensSystem &L, const PtVertex &v, const SensorDS &sds, int bsdf_id, float3 weight,
                          unsigned &n_rays, unsigned &n_splat, float (*g_loc)[kSlots]) {
    float3 so, sd; float smax;
    if (!connect_prepare(v, sds, weight, so, sd, smax)) return;
    unsigned work = 0;
    const RayHit sh = scene_intersect(A.sc, so, sd, smax, true, work);
    n_rays++;
    if (sh.tri >= 0) return;
    const float3 value = connect_value(A, v, bsdf_id, weight, sd);
    if (!kAdj) splat(A, sds.u, sds.v, value * (sds.weight * A.sample_scale), n_splat);
    else if (value.x != 0.f || value.y != 0.f || value.z != 0.f) {
        // the vertex, its BSDF value and the visibility do not depend on the lens (the aperture point's dependence on the front
        // vertex is detached, like the masks): what remains is the spot path's connection adjoint for this reference point
        if (lens_connection_adjoint(L, A.film, strict::v3(sds.ref.x, sds.ref.y, sds.ref.z), sds.a, sds.b, 0.f, value.x, value.y, value.z,
                                    A.sample_scale, 0.f, A.grad_accum, g_loc)) n_splat++;
    }
}

LRT_D float3 uniform_sphere(float sx, float sy) {      // warp.h:250-255
    const float z = strict::fnmadd(2.f, sy, 1.f), r = strict::sqrt(fmaxf(strict::fnmadd(z, z, 1.f), 0.f));
    float sn, cs; sincos_drjit(strict::mul(strict::mul(2.f, kPi), sx), sn, cs);
    return make_float3(strict::mul(r, cs), strict::mul(r, sn), z);
}

LRT_D float3 cosine_hemisphere(float sx, float sy) {   // warp.h:241-252, the same (strict) disk warp as the aperture
    float px, py; square_to_disk_concentric(sx, sy, px, py);
    return make_float3(px, py, sqrtf(fmaxf(1.f - (px * px + py * py), 0.f)));
}

// Scene::sample_emitter (scene.cpp:227-250): uniform choice, the sample is reused
LRT_D int pick_emitter(const SceneView &sc, float &sample, float &weight) {
    const int count = sc.n_emitters;
    weight = 1.f;
    if (count < 2) return 0;
    const float cf = (float) count, scaled = sample * cf;
    const int index = min((int) scaled, count - 1);
    weight = cf; sample = scaled - (float) index;
    return index;
}

// Emitter::sample_position for area lights (area.cpp:210-245): point, normal, weight = 1 / pdf
LRT_D void emitter_sample_position(const SceneView &sc, const float4 *ep, float sx, float sy, float3 &p, float3 &n, float &weight) {
    const float4 e1 = __ldg(ep + 1), e4 = __ldg(ep + 4);
    if (e4.w == 1.f) mesh_emitter_sample_position(sc, e1, sx, sy, p, n);
    else if (e4.w == 3.f) {      // Disk::sample_position (disk.cpp:173-190): concentric disk warp, then to_world
        const float4 e2 = __ldg(ep + 2), e3 = __ldg(ep + 3);
        float qx, qy; square_to_disk_concentric(sx, sy, qx, qy);
        p = make_float3(e1.x, e1.y, e1.z) + make_float3(e2.x, e2.y, e2.z) * qx + make_float3(e3.x, e3.y, e3.z) * qy;
        n = make_float3(e4.x, e4.y, e4.z);
    } else {
        const float4 e2 = __ldg(ep + 2), e3 = __ldg(ep + 3);
        const float ux = sx * 2.f - 1.f, uy = sy * 2.f - 1.f;
        p = make_float3(e1.x, e1.y, e1.z) + make_float3(e2.x, e2.y, e2.z) * ux + make_float3(e3.x, e3.y, e3.z) * uy;
        n = make_float3(e4.x, e4.y, e4.z);
    }
    weight = e1.w > 0.f ? 1.f / e1.w : 0.f;
}

// ptracer.cpp:96-167
// (up to the connection itself: the vertex on the emitter, the sensor's direction sample for it, the weight to connect with)
// part A: every draw and the vertex (v.wi is the sensor sample's direction: set by the caller); (s3x, s3y) = the aperture sample
// false: a delta emitter (`directionalarea`) — nothing to connect, its draws are done
LRT_D bool visible_emitter_vertex_a(const PtArgs &A, Pcg32 &rng, PtVertex &v, float3 &weight, float &s3x_out, float &s3y_out) {
    if (A.shutter_open) (void) rng.next_f32();
    float idx_sample = rng.next_f32(), idx_weight;
    const int index = pick_emitter(A.sc, idx_sample, idx_weight);
    const float4 *ep = A.sc.emitters + 5 * (size_t) index;
    const float4 e0 = __ldg(ep), e1 = __ldg(ep + 1), e4 = __ldg(ep + 4);
    if (__ldg(ep + 3).w == 1.f) {
        // EmitterFlags::Delta (ptracer.cpp:111): the emitter is not connected and the function's MASKED draws — position sample,
        // wavelength sample — do not advance the generator; the sensor's 2D sample (unmasked, :155) does
        (void) rng.next_f32(); (void) rng.next_f32();
        return false;
    }
    const float s2x = rng.next_f32(), s2y = rng.next_f32();
    float ew;
    if (e4.w == 2.f) {
        // a direction towards the environment as seen from the sensor's position, converted to area measure (:117-141)
        const float3 ref_p = make_float3(A.to_world[3], A.to_world[7], A.to_world[11]);
        const float3 d = uniform_sphere(s2x, s2y);
        const float3 rel = ref_p - make_float3(e1.x, e1.y, e1.z);
        const float dist = 2.f * fmaxf(e1.w, sqrtf(dot(rel, rel)));
        ew = (4.f * kPi) * (dist * dist);
        v.p = ref_p + d * dist; v.n = -d; v.shape = -1; v.uv = make_float2(s2x, s2y);
    } else {
        emitter_sample_position(A.sc, ep, s2x, s2y, v.p, v.n, ew);
        v.shape = (int) e0.w; v.uv = make_float2(0, 0);
    }
    v.sn = v.n; coordinate_system(v.n, v.s, v.t);
    const float s3x = rng.next_f32(), s3y = rng.next_f32();
    s3x_out = s3x; s3y_out = s3y;
    (void) rng.next_f32();                                       // the wavelength sample (unused in RGB variants)
    weight = make_float3(e0.x, e0.y, e0.z) * (ew * idx_weight);
    return true;
}

LRT_D bool visible_emitter_vertex(const PtArgs &A, const LensSystem &L, Pcg32 &rng, PtVertex &v, SensorDS &sds, float3 &weight) {
    float a, b;
    if (!visible_emitter_vertex_a(A, rng, v, weight, a, b)) return false;
    sds = sensor_sample_direction(A, L, v.p, a, b);
    v.wi = sds.d;
    return true;
}
template <bool kAdj>
LRT_D void sample_visible_emitters(const PtArgs &A, const LensSystem &L, Pcg32 &rng, unsigned &n_rays, unsigned &n_splat, float (*g_loc)[kSlots]) {
    PtVertex v; SensorDS sds; float3 weight;
    if (!visible_emitter_vertex(A, L, rng, v, sds, weight)) return;
    connect_sensor<kAdj>(A, L, v, sds, -1, weight, n_rays, n_splat, g_loc);
}

// ptracer.cpp:170-186 + scene.cpp:260-296 + area.cpp:92-115 / constant.cpp:93-111
LRT_D void prepare_ray(const PtArgs &A, Pcg32 &rng, float3 &ro, float3 &rd, float3 &throughput) {
    if (A.shutter_open) (void) rng.next_f32();
    float sample1 = rng.next_f32(), idx_weight;
    const float s2x = rng.next_f32(), s2y = rng.next_f32(), s3x = rng.next_f32(), s3y = rng.next_f32();
    const int index = pick_emitter(A.sc, sample1, idx_weight);
    const float4 *ep = A.sc.emitters + 5 * (size_t) index;
    const float4 e0 = __ldg(ep), e1 = __ldg(ep + 1), e4 = __ldg(ep + 4);
    const float3 radiance = make_float3(e0.x, e0.y, e0.z);
    if (e4.w == 2.f) {
        const float3 v0 = uniform_sphere(s2x, s2y);
        ro = make_float3(e1.x, e1.y, e1.z) + v0 * e1.w;
        const float3 v1 = cosine_hemisphere(s3x, s3y);
        float3 s, t; coordinate_system(-v0, s, t);
        rd = s * v1.x + t * v1.y + (-v0) * v1.z;
        throughput = radiance * ((4.f * kPi * e1.w * e1.w) * kPi);
    } else if (__ldg(ep + 3).w == 1.f) {
        // DirectionalArea::sample_ray (directionalarea.cpp:86-110): a point on the shape, direction = its normal, weight = area x radiance
        float3 p, n; float w; emitter_sample_position(A.sc, ep, s2x, s2y, p, n, w);
        rd = n; ro = offset_p(p, n, rd);
        throughput = radiance * w;
    } else {
        float3 p, n; float w; emitter_sample_position(A.sc, ep, s2x, s2y, p, n, w);
        float3 s, t; coordinate_system(n, s, t);
        const float3 local = cosine_hemisphere(s3x, s3y);
        rd = s * local.x + t * local.y + n * local.z;
        ro = offset_p(p, n, rd);
        throughput = radiance * (w * kPi);
    }
    throughput *= idx_weight;
}

// ptracer.cpp:205-283
template <bool kAdj>
LRT_D void trace_light_ray(const PtArgs &A, const LensSystem &L, Pcg32 &rng, float3 ro, float3 rd, float3 throughput,
                           unsigned &n_rays, unsigned &n_splat, float (*g_loc)[kSlots]) {
    float eta = 1.f; int depth = 1;
    bool active = throughput.x != 0.f || throughput.y != 0.f || throughput.z != 0.f;
    if (!active) return;
    unsigned work = 0;
    RayHit h = scene_intersect(A.sc, ro, rd, kLargest, false, work); n_rays++;
    active = h.tri >= 0 && depth < A.max_depth;
    while (active) {
        const Surface si = make_surface(A.sc, h, ro, rd);
        const int bsdf_id = __ldg(A.sc.shape_bsdf + si.shape);
        const float ssx = rng.next_f32(), ssy = rng.next_f32();
        const SensorDS sds = sensor_sample_direction(A, L, si.p, ssx, ssy);
        PtVertex v; v.p = si.p; v.n = si.n; v.s = si.s; v.t = si.t; v.sn = si.sn; v.wi = si.wi; v.uv = si.uv; v.shape = si.shape;
        connect_sensor<kAdj>(A, L, v, sds, bsdf_id, throughput, n_rays, n_splat, g_loc);
        // bsdf->sample(ctx, si, next_1d(), next_2d()): the 2D sample is drawn first (see the header)
        const float s2x = rng.next_f32(), s2y = rng.next_f32(), s1 = rng.next_f32();
        BsdfResult bs = bsdf_eval_pdf_sample(A.sc, bsdf_id, si.uv, si.wi, make_float3(0, 0, 1), s1, s2x, s2y);
        if (bs.transmitted) {     // TransportMode::Importance: no eta^2 factor on a refraction (dielectric.cpp:363-368)
            const float4 b2 = __ldg(A.sc.bsdfs + 4 * (size_t) bsdf_id + 2), b3 = __ldg(A.sc.bsdfs + 4 * (size_t) bsdf_id + 3);
            bs.weight = make_float3(b2.w, b3.x, b3.y);
        }
        const float3 wo_w = si.to_world(bs.wo);
        const float wi_dot_geo_n = dot(si.n, -rd), wo_dot_geo_n = dot(si.n, wo_w);
        active = (wi_dot_geo_n * si.wi.z > 0.f) && (wo_dot_geo_n * bs.wo.z > 0.f);
        const float correction = fabsf((si.wi.z * wo_dot_geo_n) / (bs.wo.z * wi_dot_geo_n));
        throughput *= bs.weight * correction;
        eta *= bs.seta;
        active = active && (throughput.x != 0.f || throughput.y != 0.f || throughput.z != 0.f);
        if (!active) break;
        ro = offset_p(si.p, si.n, wo_w); rd = wo_w;
        h = scene_intersect(A.sc, ro, rd, kLargest, false, work); n_rays++;
        depth++;
        active = depth < A.max_depth && h.tri >= 0;
        const bool use_rr = depth > A.rr_depth;
        if (use_rr || !A.scalar) {
            const float q = fminf(max3(throughput) * eta * eta, .95f);
            const float r = active ? rng.next_f32() : 0.f;          // a masked draw: inactive lanes keep their state
            if (use_rr) { active = active && r < q; throughput *= 1.f / q; }
        }
    }
}

// ptracer.cpp:77-90
template <bool kAdj>
LRT_D void ptrace_sample(const PtArgs &A, const LensSystem &L, Pcg32 &rng, unsigned &n_rays, unsigned &n_splat, float (*g_loc)[kSlots]) {
    if (A.max_depth != 0 && !A.hide_emitters) sample_visible_emitters<kAdj>(A, L, rng, n_rays, n_splat, g_loc);
    float3 ro, rd, throughput;
    prepare_ray(A, rng, ro, rd, throughput);
    trace_light_ray<kAdj>(A, L, rng, ro, rd, throughput, n_rays, n_splat, g_loc);
}

template <bool kAdj>
__global__ void __launch_bounds__(128) k_ptrace(PtArgs A) {
    __shared__ LensSystem L;
    __shared__ float s_grad[kAdj ? kMaxInterfaces * kSlots : 1];
    float g_loc[kAdj ? kMaxInterfaces : 1][kSlots];
    if (kAdj) {
        for (int i = threadIdx.x; i < kMaxInterfaces * kSlots; i += blockDim.x) s_grad[i] = 0.f;
        for (int i = 0; i < kMaxInterfaces; ++i) for (int k = 0; k < kSlots; ++k) g_loc[i][k] = 0.f;
    }
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.lens);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    unsigned n_paths = 0, n_rays = 0, n_splat = 0;
    const long long tid = blockIdx.x * (long long) blockDim.x + threadIdx.x, stride = (long long) gridDim.x * blockDim.x;
    if (!A.scalar) {
        // sample idx = shard_index + k * shard_count: identical streams for any number of GPUs
        for (long long k = tid; ; k += stride) {
            const long long idx = (long long) A.shard_index + k * A.shard_count;
            if (idx >= A.n_total) break;
            Pcg32 rng; rng.seed_sample(A.seed, (uint32_t) idx);
            ptrace_sample<kAdj>(A, L, rng, n_rays, n_splat, g_loc); n_paths++;
        }
    } else {
        for (long long r = tid; r < A.n_ranges; r += stride) {
            const long long begin = r * A.grain, end = min(begin + A.grain, A.n_total);
            Pcg32 rng; rng.seed((uint64_t) (A.seed + (uint32_t) r), 0xda3e39cb94b95bdbULL);
            for (long long i = begin; i < end; ++i) { ptrace_sample<kAdj>(A, L, rng, n_rays, n_splat, g_loc); n_paths++; }
        }
    }
    if (kAdj) {     // per-thread slots -> warp -> CTA -> one global atomic per slot (as the spot adjoint)
        const int n_if = L.n;
        for (int i = 0; i < n_if; ++i)
            for (int k = 0; k < kSlots; ++k) {
                float v = g_loc[i][k];
                for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
                if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(&s_grad[i * kSlots + k], v);
            }
        __syncthreads();
        for (int i = threadIdx.x; i < n_if * kSlots; i += blockDim.x) {
            const float v = s_grad[i];
            if (v == 0.f) continue;
            const int r = i / kSlots, k = i - r * kSlots;
            const int col = k == 0 ? G_Z : (k == 1 ? G_R : (k == 2 ? G_K : (k < 10 ? G_A0 + (k - 3) : (k == 10 ? G_IORL : G_IORR))));
            atomicAdd(A.grad_rows + r * kGradRow + col, v);
        }
    }
    if (A.stats) {
        unsigned long long c[3] = { n_paths, n_splat, n_rays };
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xffffffffu, c[k], o);
            if ((threadIdx.x & 31) == 0 && c[k]) atomicAdd(A.stats + k, c[k]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// The default kernel (per-sample streams; forward renders and, kAdj, the lens adjoint): the same light paths with the render
// megakernel's execution scheme (lrt_render.cu) instead of one thread per path run to completion. Persistent warps; every lane
// carries one light path as a small state machine — phase 1: a path segment (closest hit), phase 2: the shadow ray of a
// connection — and all lanes advance together through ONE traversal site in warp-uniform bursts (bvh_burst; by default a burst
// runs until all its lanes are done: LRT_PT_EXIT_DROP = 32 measured best, the shading step below being the expensive part).
// Lanes whose ray is finished do their shading step — ONE site for the sensor's direction sample (a strict lens trace, half of
// all instructions) shared by the emitter vertices of new paths and the scene vertices of running ones, then the connection's
// shadow ray or the BSDF sample — and a lane whose path has ended takes the next sample index of the warp's chunk at once
// (profiles/r02w_ptrace_lines.txt: 5.4 lanes of traversal in the per-thread kernel; r02z_ptrace_queue_lines.txt: this one).
// Every lane draws its samples in the reference's order: the image equals k_ptrace's up to the order of the atomics.
constexpr int kPtChunk = 64;          // sample indices handed to a warp per fetch

template <bool kAdj>
__global__ void __launch_bounds__(128, 4) k_ptrace_queue(PtArgs A, unsigned long long *counter, long long n_my, int exit_drop, int leaf_min) {
    __shared__ LensSystem L;
    __shared__ float s_grad[kAdj ? kMaxInterfaces * kSlots : 1];
    float g_loc[kAdj ? kMaxInterfaces : 1][kSlots];       // kAdj: this thread's share of d(loss)/d(interface table)
    if (kAdj) {
        for (int i = threadIdx.x; i < kMaxInterfaces * kSlots; i += blockDim.x) s_grad[i] = 0.f;
        for (int i = 0; i < kMaxInterfaces; ++i) for (int q = 0; q < kSlots; ++q) g_loc[i][q] = 0.f;
    }
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.lens);
        uint4 *dst = reinterpret_cast<uint4 *>(&L);
        for (int i = threadIdx.x; i < (int) (sizeof(LensSystem) / 16); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned n_paths = 0, n_rays = 0, n_splat = 0;
    uint2 st[kTravStack];
    // ---- per-lane path state
    int phase = 0;                     // 0 idle, 1 path segment, 2 shadow ray of a connection
    bool stage0 = false, first_ray = false;   // stage0: the connection in flight is the emitter vertex's (next: the emitter ray)
    Pcg32 rng; rng.state = 0; rng.inc = 1;
    float3 ro = make_float3(0, 0, 0), rd = make_float3(0, 0, 1), throughput = make_float3(0, 0, 0);
    float3 seg_o = ro, seg_d = rd;     // the segment that produced the current vertex (its surface is rebuilt after the shadow ray)
    RayHit seg; seg.t = 0.f; seg.u = seg.v = 0.f; seg.tri = -1;
    float eta = 1.f; int depth = 1;
    float pu = 0.f, pv = 0.f; float3 pval = make_float3(0, 0, 0);      // the connection waiting for its shadow ray
    float3 pref = make_float3(0, 0, 0);                                  // kAdj: its camera-space reference point; (pu, pv) = the aperture sample
    Trav T; trav_init(T, 0.f); T.cur = kTravDone;
    // analytic spheres (outside the BVH) are tested when a ray starts, as in the render kernel: the traversal then runs with the
    // sphere's distance as its limit, a shadow ray that a sphere blocks skips it; the sphere's index + 1 rides in T.hit.u
    auto start_ray = [&](float maxt, bool any_hit) {
        int sph = -1;
        if (A.sc.n_spheres > 0) {
            float t = maxt;
            sph = spheres_intersect(A.sc, ro, rd, maxt, any_hit, t);
            if (sph >= 0) maxt = t;
        }
        trav_init(T, maxt);
        if (sph >= 0) { T.hit.u = __int_as_float(sph + 1); if (any_hit) T.cur = kTravDone; }
    };
    long long next = 0, end = 0;       // the warp's chunk of sample numbers k (idx = shard_index + k * shard_count), warp-uniform
    bool more = true;
    while (true) {
        // ---- A. idle lanes take the next sample indices (a new chunk whenever the current one is used up)
        const unsigned im = __ballot_sync(0xffffffffu, phase == 0);
        if (im != 0u && next >= end && more) {
            long long c = 0;
            if (lane == 0) c = (long long) atomicAdd(counter, 1ULL);
            c = __shfl_sync(0xffffffffu, c, 0);
            next = c * kPtChunk; end = min(next + kPtChunk, n_my);
            if (next >= n_my) { more = false; next = end = 0; }
        }
        const int take = (int) min((long long) __popc(im), end - next), rank = __popc(im & lt_mask);
        bool fresh = false;
        if (phase == 0 && rank < take) {
            const long long idx = (long long) A.shard_index + (next + rank) * A.shard_count;
            rng.seed_sample(A.seed, (uint32_t) idx);
            n_paths++; fresh = true;
        }
        next += take;
        if (__ballot_sync(0xffffffffu, phase != 0 || fresh) == 0u) {
            if (!more) break;
            continue;
        }
        // ---- B. the one traversal site: segments and shadow rays of different lanes advance together
        unsigned work = 0;
        bvh_burst(A.sc.bvh, ro, rd, T, st, phase != 0, phase == 2, exit_drop, leaf_min, work);
        // ---- C. lanes whose ray is finished (or that have just taken a sample) do their shading step
        if (fresh || (phase != 0 && T.cur == kTravDone)) {
            bool need_next = false, have_si = false, want_sds = false, on_emitter = false;
            Surface si; int bsdf_id = -1;
            PtVertex v; float3 w_conn = make_float3(0, 0, 0); float sa = 0.f, sb = 0.f;
            if (fresh) {
                stage0 = true; first_ray = true; depth = 1; eta = 1.f;
                if (A.max_depth != 0 && !A.hide_emitters && visible_emitter_vertex_a(A, rng, v, w_conn, sa, sb)) want_sds = on_emitter = true;
                else need_next = true;
            } else if (phase == 2) {
                n_rays++;
                if (T.hit.tri < 0 && T.hit.u == 0.f) {                                     // (u != 0 without a triangle: a sphere is in the way)
                    if (!kAdj) splat(A, pu, pv, pval, n_splat);
                    else if (pval.x != 0.f || pval.y != 0.f || pval.z != 0.f) {
                        // the connection's lens adjoint (see connect_sensor<true>): the spot path's adjoint for this reference point
                        if (lens_connection_adjoint(L, A.film, strict::v3(pref.x, pref.y, pref.z), pu, pv, 0.f, pval.x, pval.y, pval.z,
                                                    A.sample_scale, 0.f, A.grad_accum, g_loc)) n_splat++;
                    }
                }
                need_next = true;
            } else {
                RayHit h = T.hit;
                if (h.tri < 0 && h.u != 0.f) { h.tri = A.sc.n_tris + __float_as_int(h.u) - 1; h.t = T.tmax; h.u = 0.f; }     // the sphere is the closest hit
                n_rays++;
                if (!first_ray) depth++;
                bool active = depth < A.max_depth && h.tri >= 0;
                if (!first_ray) {
                    const bool use_rr = depth > A.rr_depth;
                    const float q = fminf(max3(throughput) * eta * eta, .95f);
                    const float r = active ? rng.next_f32() : 0.f;          // a masked draw: inactive lanes keep their state
                    if (use_rr) { active = active && r < q; throughput *= 1.f / q; }
                }
                first_ray = false;
                if (!active) phase = 0;
                else {
                    seg = h; seg_o = ro; seg_d = rd;
                    si = make_surface(A.sc, h, ro, rd); have_si = true;
                    bsdf_id = __ldg(A.sc.shape_bsdf + si.shape);
                    sa = rng.next_f32(); sb = rng.next_f32();
                    v.p = si.p; v.n = si.n; v.s = si.s; v.t = si.t; v.sn = si.sn; v.wi = si.wi; v.uv = si.uv; v.shape = si.shape;
                    w_conn = throughput; want_sds = true;
                }
            }
            // the ONE site of the sensor's direction sample (a lens trace front -> rear in strict arithmetic, the most expensive
            // step of a vertex): emitter vertices of new paths and scene vertices of running paths go through it together
            if (want_sds) {
                const SensorDS sds = sensor_sample_direction(A, L, v.p, sa, sb);
                if (on_emitter) v.wi = sds.d;
                float3 so, sd; float smax;
                if (connect_prepare(v, sds, w_conn, so, sd, smax)) {
                    pval = connect_value(A, v, bsdf_id, w_conn, sd);
                    if (!kAdj) { pval *= sds.weight * A.sample_scale; pu = sds.u; pv = sds.v; }
                    else { pref = sds.ref; pu = sds.a; pv = sds.b; }
                    ro = so; rd = sd; start_ray(smax, true); phase = 2;
                } else need_next = true;
            }
            if (need_next) {
                if (stage0) {            // the emitter ray (prepare_ray)
                    stage0 = false;
                    prepare_ray(A, rng, ro, rd, throughput);
                    if (throughput.x != 0.f || throughput.y != 0.f || throughput.z != 0.f) { start_ray(kLargest, false); phase = 1; }
                    else phase = 0;
                } else {                 // sample the BSDF at the current vertex
                    if (!have_si) { si = make_surface(A.sc, seg, seg_o, seg_d); bsdf_id = __ldg(A.sc.shape_bsdf + si.shape); }
                    // bsdf->sample(ctx, si, next_1d(), next_2d()): the 2D sample is drawn first (see the header)
                    const float s2x = rng.next_f32(), s2y = rng.next_f32(), s1 = rng.next_f32();
                    BsdfResult bs = bsdf_eval_pdf_sample(A.sc, bsdf_id, si.uv, si.wi, make_float3(0, 0, 1), s1, s2x, s2y);
                    if (bs.transmitted) {     // TransportMode::Importance: no eta^2 factor on a refraction (dielectric.cpp:363-368)
                        const float4 b2 = __ldg(A.sc.bsdfs + 4 * (size_t) bsdf_id + 2), b3 = __ldg(A.sc.bsdfs + 4 * (size_t) bsdf_id + 3);
                        bs.weight = make_float3(b2.w, b3.x, b3.y);
                    }
                    const float3 wo_w = si.to_world(bs.wo);
                    const float wi_dot_geo_n = dot(si.n, -seg_d), wo_dot_geo_n = dot(si.n, wo_w);
                    bool active = (wi_dot_geo_n * si.wi.z > 0.f) && (wo_dot_geo_n * bs.wo.z > 0.f);
                    const float correction = fabsf((si.wi.z * wo_dot_geo_n) / (bs.wo.z * wi_dot_geo_n));
                    throughput *= bs.weight * correction;
                    eta *= bs.seta;
                    active = active && (throughput.x != 0.f || throughput.y != 0.f || throughput.z != 0.f);
                    if (active) { ro = offset_p(si.p, si.n, wo_w); rd = wo_w; start_ray(kLargest, false); phase = 1; }
                    else phase = 0;
                }
            }
        }
        __syncwarp();
    }
    if (kAdj) {     // per-thread slots -> warp -> CTA -> one global atomic per slot (as k_ptrace<true> and the spot adjoint)
        const int n_if = L.n;
        for (int i = 0; i < n_if; ++i)
            for (int q = 0; q < kSlots; ++q) {
                float v = g_loc[i][q];
                for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
                if (lane == 0 && v != 0.f) atomicAdd(&s_grad[i * kSlots + q], v);
            }
        __syncthreads();
        for (int i = threadIdx.x; i < n_if * kSlots; i += blockDim.x) {
            const float v = s_grad[i];
            if (v == 0.f) continue;
            const int r = i / kSlots, q = i - r * kSlots;
            const int col = q == 0 ? G_Z : (q == 1 ? G_R : (q == 2 ? G_K : (q < 10 ? G_A0 + (q - 3) : (q == 10 ? G_IORL : G_IORR))));
            atomicAdd(A.grad_rows + r * kGradRow + col, v);
        }
    }
    if (A.stats) {
        unsigned long long c[3] = { n_paths, n_splat, n_rays };
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xffffffffu, c[k], o);
            if (lane == 0 && c[k]) atomicAdd(A.stats + k, c[k]);
        }
    }
}

}  // namespace lrt

using namespace lrt;
extern "C" {

static int ptrace_launch(const lrt_scene *sc, const lrt_sensor *sensor_, const lrt_film_desc *film, const lrt_ptrace_params *pp, float *accum_dev,
                         const float *grad_accum_dev, float *grad_rows_dev, void *stream_) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && sensor_ && film && pp && (accum_dev || (grad_accum_dev && grad_rows_dev)), "null argument");
    LRT_REQUIRE(!grad_accum_dev || pp->rng_mode == 0, "the lens adjoint follows the per-sample streams (rng_mode 0)");
    lrt_sensor *sensor = const_cast<lrt_sensor *>(sensor_);
    cudaStream_t stream = (cudaStream_t) stream_;
    LRT_REQUIRE(sensor->host.kind == 0 || !grad_accum_dev, "the lens adjoint needs a realistic lens sensor");
    LRT_REQUIRE(film->channels == 4 && film->normalize == 1 && film->coalesce == 0,
                "lrt_ptrace needs the particle tracer's film: 4 channels, normalize = 1, coalesce = 0 (integrator.cpp:706-717)");
    LRT_REQUIRE(pp->spp > 0 && pp->max_depth >= -1 && pp->rr_depth > 0, "bad parameters (spp > 0, max_depth >= -1, rr_depth > 0)");
    LRT_REQUIRE(pp->shard_count >= 1 && pp->shard_index >= 0 && pp->shard_index < pp->shard_count, "bad shard");
    LRT_REQUIRE(pp->rng_mode == 0 || pp->rng_mode == 1, "rng_mode must be 0 (per-sample streams) or 1 (scalar convention)");
    LRT_REQUIRE(pp->rng_mode == 0 || (pp->shard_count == 1 && pp->n_threads >= 1), "rng_mode 1 is a single-GPU parity mode and needs n_threads >= 1");
    if (sc->n_emitters == 0) return LRT_OK;                       // "no emitters found, returning black image"
    PtArgs A; memset(&A, 0, sizeof A);
    int rc = film_params_from_desc(film, A.film); if (rc) return rc;
    rc = sensor->upload(stream); if (rc) return rc;
    A.sc = sc->view; A.lens = sensor->dev;
    for (int i = 0; i < 12; ++i) A.to_world[i] = sensor->host.to_world[i];
    bool given = false;
    for (int i = 0; i < 12; ++i) given = given || pp->to_camera[i] != 0.f;
    if (given) for (int i = 0; i < 12; ++i) A.to_camera[i] = pp->to_camera[i];
    else {    // numerical inverse of the affine matrix, in double, rounded once
        const float *T = A.to_world;
        const double a = T[0], b = T[1], c = T[2], d = T[4], e = T[5], f = T[6], g = T[8], h = T[9], k = T[10];
        const double det = a * (e * k - f * h) - b * (d * k - f * g) + c * (d * h - e * g);
        LRT_REQUIRE(det != 0.0, "the sensor's to_world matrix is singular");
        const double R[9] = { (e * k - f * h) / det, (c * h - b * k) / det, (b * f - c * e) / det, (f * g - d * k) / det, (a * k - c * g) / det,
                              (c * d - a * f) / det, (d * h - e * g) / det, (b * g - a * h) / det, (a * e - b * d) / det };
        for (int r = 0; r < 3; ++r) {
            for (int q = 0; q < 3; ++q) A.to_camera[4 * r + q] = (float) R[3 * r + q];
            A.to_camera[4 * r + 3] = (float) -(R[3 * r] * T[3] + R[3 * r + 1] * T[7] + R[3 * r + 2] * T[11]);
        }
    }
    const lrt_film_geom &g = sensor->film;
    if (sensor->host.kind == 1) {
        // perspective sensor: camera -> sample (sensor.h:228-265: scale * translate * scale * translate * perspective, composed in
        // double and rounded once), the image rectangle at z = 1 and its normalisation (perspective.cpp:186-194)
        A.persp = 1; A.near_clip = sensor->host.near_clip; A.far_clip = sensor->host.far_clip;
        const double W = g.width, H = g.height, aspect = W / H, relx = g.crop_w / W, rely = g.crop_h / H, offx = g.crop_x / W, offy = g.crop_y / H;
        const double n_ = A.near_clip, f_ = A.far_clip, recip = 1.0 / (f_ - n_), cot = 1.0 / std::tan(0.5 * (double) sensor->persp_fov * 3.14159265358979323846 / 180.0);
        auto mul4 = [](const double a[16], const double b[16], double out[16]) {
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { double v = 0; for (int k = 0; k < 4; ++k) v += a[4 * i + k] * b[4 * k + j]; out[4 * i + j] = v; }
        };
        const double P[16] = { cot, 0, 0, 0, 0, cot, 0, 0, 0, 0, f_ * recip, -n_ * f_ * recip, 0, 0, 1, 0 };
        const double T2[16] = { 1, 0, 0, -1, 0, 1, 0, -1.0 / aspect, 0, 0, 1, 0, 0, 0, 0, 1 };
        const double S2[16] = { -0.5, 0, 0, 0, 0, -0.5 * aspect, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
        const double T1[16] = { 1, 0, 0, -offx, 0, 1, 0, -offy, 0, 0, 1, 0, 0, 0, 0, 1 };
        const double S1[16] = { 1.0 / relx, 0, 0, 0, 0, 1.0 / rely, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
        double a[16], b[16], c[16], m[16];
        mul4(T2, P, a); mul4(S2, a, b); mul4(T1, b, c); mul4(S1, c, m);
        for (int i = 0; i < 16; ++i) A.c2s[i] = (float) m[i];
        const float *q = sensor->host.persp_s2c;
        const float x0 = q[3] / q[11], y0 = q[7] / q[11], x1 = (q[0] + q[3]) / q[11], y1 = (q[5] + q[7]) / q[11];
        A.rect_min[0] = std::min(x0, x1); A.rect_max[0] = std::max(x0, x1); A.rect_min[1] = std::min(y0, y1); A.rect_max[1] = std::max(y0, y1);
        A.persp_norm = 1.f / ((A.rect_max[0] - A.rect_min[0]) * (A.rect_max[1] - A.rect_min[1]));
        A.persp_res[0] = (float) g.crop_w; A.persp_res[1] = (float) g.crop_h;
    }
    const long long film_pixels = (long long) g.width * g.height;
    A.n_total = film_pixels * pp->spp;                            // integrator.cpp:573-574: the FULL film size x spp
    LRT_REQUIRE(A.n_total <= 0xffffffffll, "width * height * spp must stay below 2^32 (split the render into passes)");
    A.sample_scale = (float) ((long long) film->size_x * film->size_y) / (float) ((uint32_t) pp->spp * (uint32_t) film_pixels);
    A.scalar = pp->rng_mode; A.max_depth = pp->max_depth < 0 ? 0x7fffffff : pp->max_depth; A.rr_depth = pp->rr_depth;
    A.hide_emitters = pp->hide_emitters; A.shutter_open = sensor->host.shutter_open_time > 0.f;
    A.shard_index = pp->shard_index; A.shard_count = pp->shard_count;
    A.film.discretized = A.scalar;
    long long work_items;
    if (A.scalar) {
        A.grain = std::max<long long>(A.n_total / (4ll * pp->n_threads), 1);
        A.n_ranges = (A.n_total + A.grain - 1) / A.grain;
        A.seed = pp->base_seed + pp->seed * (uint32_t) ((uint32_t) A.n_total / (uint32_t) A.grain);   // integrator.cpp:621 + sampler.cpp:131
        work_items = A.n_ranges;
    } else {
        A.seed = pp->base_seed + pp->seed;
        work_items = (A.n_total - pp->shard_index + pp->shard_count - 1) / pp->shard_count;
    }
    A.accum = accum_dev; A.grad_accum = grad_accum_dev; A.grad_rows = grad_rows_dev; A.stats = sc->d_counter + kCounterRing;
    LRT_CUDA(cudaMemsetAsync(A.stats, 0, 8 * sizeof(unsigned long long), stream));
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long grid = (work_items + 127) / 128; grid = std::max<long long>(1, std::min<long long>(grid, (long long) sms * 16));
    // the default convention takes the queue kernel (LRT_PTRACE_SIMPLE=1: the per-thread one); the scalar parity mode the per-thread one
    static const bool simple = getenv("LRT_PTRACE_SIMPLE") != nullptr && atoi(getenv("LRT_PTRACE_SIMPLE")) != 0;
    if (A.scalar || simple) {
        if (grad_accum_dev) k_ptrace<true><<<(int) grid, 128, 0, stream>>>(A);
        else k_ptrace<false><<<(int) grid, 128, 0, stream>>>(A);
    } else {
        const char *e = getenv("LRT_PT_EXIT_DROP"); int exit_drop = e ? atoi(e) : 32;
        exit_drop = std::max(1, std::min(32, exit_drop));
        e = getenv("LRT_PT_LEAF_MIN"); const int leaf_min = std::max(1, e ? atoi(e) : 6);
        const long long chunks = (work_items + kPtChunk - 1) / kPtChunk;
        const int qgrid = (int) std::max<long long>(1, std::min<long long>((chunks + 3) / 4, (long long) sms * 4));
        // the chunk counter: a slot of the scene's per-launch ring (as lrt_render), so that particle-tracer launches of one
        // scene on different streams do not share their work distribution
        unsigned long long *ctr = sc->d_counter + (sc->launch_seq.fetch_add(1, std::memory_order_relaxed) % kCounterRing);
        LRT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(unsigned long long), stream));
        if (grad_accum_dev) k_ptrace_queue<true><<<qgrid, 128, 0, stream>>>(A, ctr, work_items, exit_drop, leaf_min);
        else k_ptrace_queue<false><<<qgrid, 128, 0, stream>>>(A, ctr, work_items, exit_drop, leaf_min);
    }
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_ptrace(const lrt_scene *sc, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_ptrace_params *pp, float *accum_dev,
               void *stream) {
    LRT_REQUIRE(accum_dev, "null argument");
    return ptrace_launch(sc, sensor, film, pp, accum_dev, nullptr, nullptr, stream);
}

int lrt_ptrace_backward(const lrt_scene *sc, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_ptrace_params *pp,
                        const float *grad_accum_dev, float *grad_rows16_dev, void *stream) {
    LRT_REQUIRE(grad_accum_dev && grad_rows16_dev, "null argument");
    return ptrace_launch(sc, sensor, film, pp, nullptr, grad_accum_dev, grad_rows16_dev, stream);
}

}  // extern "C"
