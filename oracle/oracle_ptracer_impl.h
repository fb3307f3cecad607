// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). Implementation fragment, included at the end of oracle_path.cpp (it uses
// that file's scene, intersection, BSDF and emitter restatements). The particle tracer:
//   ParticleTracerIntegrator::sample / sample_visible_emitters / prepare_ray / trace_light_ray / connect_sensor
//                                                                src/integrators/ptracer.cpp:77-371
//   AdjointIntegrator::render (scalar branch: ranges of `grain_size` samples, one sampler seed per range; JIT branch: one
//   PCG32 stream per sample)                                     src/render/integrator.cpp:544-760
//   Scene::sample_emitter / sample_emitter_ray                   src/render/scene.cpp:227-250,260-296
//   AreaLight::sample_ray / sample_position                      src/emitters/area.cpp:92-115,210-245
//   ConstantBackgroundEmitter::sample_ray                        src/emitters/constant.cpp:93-111
//   RealisticLensCamera::sample_direction (world space)          src/sensors/realisticlens.cpp:2301-2394 (camera-space part:
//                                                                oracle_spot.cpp camera_sample_direction, pinned)
//   ImageBlock::put, normalize = true, weight 0                  src/render/imageblock.cpp:264-379
// Pinned against the reference's own scalar_rgb build: tests/golden/ref_ptracer.npz (tests/test_ptracer_parity.py).

namespace orc {

bool camera_sample_direction(const Camera &c, V3 ref_p, V2 sample, float *out11, float origin_offset);   // oracle_spot.cpp

struct SensorDS { bool valid = false; V2 uv { 0, 0 }; V3 p { 0, 0, 0 }, d { 0, 0, 0 }; float dist = 0, pdf = 0, weight = 0; };

struct PtCtx {
    const Scene *s; const Camera *cam;
    float to_world[12], to_camera[12];       // rows of the sensor's to_world and of its inverse
    float shutter_open_time;
    int max_depth, rr_depth; bool hide_emitters;
    float sample_scale; bool scalar;           // scalar variants splat with the discretised filter
    Filter flt; int cw, ch, cx, cy;
    // perspective sensor (cam == nullptr): camera -> sample projection (sensor.h:228-265), the image rectangle at z = 1 and
    // its normalisation (perspective.cpp:174-194), the crop resolution
    float c2s[16]; float rect_min[2], rect_max[2]; float normalization; float res[2]; float near_clip, far_clip;
};

// the perspective sensor of the next orc_ptracer_render / orc_ptracer_sample call with cam == NULL
static struct { float fov = 39.3077f, near_clip = 1e-2f, far_clip = 1e4f; } g_pt_persp;

// PerspectiveCamera::sample_direction (src/sensors/perspective.cpp:283-321) + importance (:337-386), camera space
static bool persp_sample_direction(const PtCtx &C, V3 ref_p, float out[11]) {
    if (!(ref_p.z >= C.near_clip && ref_p.z <= C.far_clip)) return false;
    const float *m = C.c2s;
    // Transform4f * Point3f: homogeneous product, then the division by w (transform.h:118-131)
    float x = fmadd(m[2], ref_p.z, fmadd(m[1], ref_p.y, fmadd(m[0], ref_p.x, m[3]))),
          y = fmadd(m[6], ref_p.z, fmadd(m[5], ref_p.y, fmadd(m[4], ref_p.x, m[7]))),
          w = fmadd(m[14], ref_p.z, fmadd(m[13], ref_p.y, fmadd(m[12], ref_p.x, m[15])));
    float iw = rcp(w);
    float u = x * iw, v = y * iw;
    if (!(u >= 0.f && u <= 1.f && v >= 0.f && v <= 1.f)) return false;
    float dist = norm(ref_p), inv_dist = rcp(dist);
    V3 local_d = ref_p * inv_dist;
    float ct = local_d.z, inv_ct = rcp(ct);
    float px = local_d.x * inv_ct, py = local_d.y * inv_ct;
    bool valid = ct > 0.f && px >= C.rect_min[0] && px <= C.rect_max[0] && py >= C.rect_min[1] && py <= C.rect_max[1];
    float importance = valid ? C.normalization * inv_ct * inv_ct * inv_ct : 0.f;
    out[0] = u * C.res[0]; out[1] = v * C.res[1];
    out[2] = out[3] = out[4] = 0.f;                 // ds.p = the pinhole (camera space origin)
    out[8] = dist; out[9] = 1.f; out[10] = importance * inv_dist * inv_dist;
    return true;
}

// Sensor::sample_direction for the realistic lens, world space in and out
static SensorDS sensor_sample_direction(const PtCtx &C, V3 p_world, V2 sample) {
    SensorDS ds;
    V3 ref_p = xf_point(C.to_camera, p_world);
    float o[11];
    if (!C.cam) { if (!persp_sample_direction(C, ref_p, o)) return ds; }
    else if (!camera_sample_direction(*C.cam, ref_p, sample, o, 0.f)) return ds;
    ds.valid = true;
    ds.uv = { o[0], o[1] };
    ds.p = xf_point(C.to_world, V3 { o[2], o[3], o[4] });
    ds.dist = o[8];
    ds.d = (ds.p - p_world) * rcp(ds.dist);
    ds.pdf = o[9]; ds.weight = o[10];
    return ds;
}

// what connect_sensor needs to know about the vertex it connects
struct PtVertex {
    V3 p, n;                 // position, geometric normal
    V3 sh_s, sh_t, sh_n;     // shading frame
    V3 wi;                   // local incident direction (surface vertices)
    V2 uv { 0, 0 };
    int shape = -1;          // -1: not on a shape (environment)
    V3 to_local(V3 v) const { return { dot(v, sh_s), dot(v, sh_t), dot(v, sh_n) }; }
    V3 to_world(V3 v) const {
        return { fmadd(sh_n.x, v.z, fmadd(sh_t.x, v.y, sh_s.x * v.x)), fmadd(sh_n.y, v.z, fmadd(sh_t.y, v.y, sh_s.y * v.x)),
                 fmadd(sh_n.z, v.z, fmadd(sh_t.z, v.y, sh_s.z * v.x)) };
    }
};

// BSDF::eval(ctx, si, wo) for the connection: only the diffuse lobe is non-zero (delta lobes evaluate to 0)
static V3 bsdf_eval(const Scene &s, const Bsdf &b, V2 uv, V3 wi, V3 wo) {
    if (b.twosided && wi.z < 0.f) { wi.z = -wi.z; wo.z = -wo.z; }
    if (b.type != 0) return { 0, 0, 0 };
    if (!(wi.z > 0.f && wo.z > 0.f)) return { 0, 0, 0 };
    return bsdf_reflectance(s, b, uv) * InvPi * wo.z;
}

// ptracer.cpp:285-371. bsdf == nullptr: a directly visible emitter (no scattering)
static void connect_sensor(const PtCtx &C, const PtVertex &v, const SensorDS &sds, const Bsdf *bsdf, V3 weight, double *block) {
    if (!(sds.valid && sds.pdf > 0.f && (weight.x != 0.f || weight.y != 0.f || weight.z != 0.f))) return;
    Ray sensor_ray = spawn_ray_to(v.p, v.n, sds.p);
    Hit h = intersect(*C.s, sensor_ray, true);
    if (h.prim >= 0 || h.rect >= 0 || h.sphere >= 0) return;
    V3 surface_weight { 1, 1, 1 };
    V3 local_d = v.to_local(sensor_ray.d);
    if (v.shape >= 0) {
        if (!bsdf) surface_weight = surface_weight * std::fmax(0.f, local_d.z);
        else {
            float wi_dot_geo_n = dot(v.n, v.to_world(v.wi)), wo_dot_geo_n = dot(v.n, sensor_ray.d);
            bool valid = (wi_dot_geo_n * v.wi.z > 0.f) && (wo_dot_geo_n * local_d.z > 0.f);
            float correction = valid ? std::fabs((v.wi.z * wo_dot_geo_n) / (local_d.z * wi_dot_geo_n)) : 0.f;
            surface_weight = surface_weight * (bsdf_eval(*C.s, *bsdf, v.uv, v.wi, local_d) * correction);
        }
    } else if (!bsdf && local_d.z <= 0.f) surface_weight = { 0, 0, 0 };
    V3 result = weight * surface_weight * C.sample_scale;
    float vals[4] = { result.x, result.y, result.z, 0.f };
    // block offset = crop offset: `sensor_ds.uv + block->offset()` then put() subtracts it again
    block_put(C.flt, C.cw, C.ch, C.cx, C.cy, 0, 4, false, true, C.scalar, sds.uv.x + (float) C.cx, sds.uv.y + (float) C.cy, vals, block);
}

// Frame3f(n): coordinate_system (frame.h:25-27)
static void frame_from_normal(PtVertex &v, V3 n) { v.sh_n = n; coordinate_system(n, v.sh_s, v.sh_t); }

// Emitter::sample_position for area lights (area.cpp:210-245): point, normal, 1 / pdf
static void emitter_sample_position(const Emitter &e, V2 sample, V3 &p, V3 &n, float &weight) {
    if (e.kind == 1) mesh_sample_position(e, sample, p, n);
    else if (e.kind == 3) {      // Disk::sample_position (disk.cpp:173-190): concentric disk warp, to_world.transform_affine
        V2 q = square_to_uniform_disk_concentric(sample);
        p = { fmadd(e.ey.x, q.y, fmadd(e.ex.x, q.x, e.c.x)), fmadd(e.ey.y, q.y, fmadd(e.ex.y, q.x, e.c.y)), fmadd(e.ey.z, q.y, fmadd(e.ex.z, q.x, e.c.z)) };
        n = e.n;
    } else {
        float sx = sample.x * 2.f - 1.f, sy = sample.y * 2.f - 1.f;
        p = { fmadd(e.ey.x, sy, fmadd(e.ex.x, sx, e.c.x)), fmadd(e.ey.y, sy, fmadd(e.ex.y, sx, e.c.y)), fmadd(e.ey.z, sy, fmadd(e.ex.z, sx, e.c.z)) };
        n = e.n;
    }
    weight = e.inv_area > 0.f ? rcp(e.inv_area) : 0.f;
}

static V3 uniform_sphere(V2 sample) {     // warp.h:250-255
    float z = fnmadd(2.f, sample.y, 1.f), r = std::sqrt(std::fmax(fnmadd(z, z, 1.f), 0.f)), sn, cs;
    sincos(2.f * Pi * sample.x, sn, cs);
    return { r * cs, r * sn, z };
}

// ptracer.cpp:96-167
static void sample_visible_emitters(const PtCtx &C, PCG32 &rng, double *block) {
    const Scene &s = *C.s;
    if (C.shutter_open_time > 0.f) (void) rng.next_float32();
    float idx_sample = rng.next_float32();
    size_t count = s.emitters.size();
    int index = 0; float idx_weight = 1.f;
    if (count > 1) {
        float cf = (float) count, scaled = idx_sample * cf;
        index = (int) std::min((uint32_t) scaled, (uint32_t) count - 1u); idx_weight = cf;
    }
    const Emitter &e = s.emitters[index];
    if (e.directional) {
        // EmitterFlags::Delta (ptracer.cpp:111): the emitter is not connected, and the MASKED draws of the function — position
        // sample, wavelength sample — do not advance the generator; the sensor's 2D sample (unmasked, :155) does
        (void) rng.next_float32(); (void) rng.next_float32();
        return;
    }
    PtVertex v; V3 emitter_weight { 0, 0, 0 };
    V2 s2; s2.x = rng.next_float32(); s2.y = rng.next_float32();
    if (e.kind == 2) {
        // 3.a: a direction towards the environment as seen from the sensor's position; converted to area measure
        V3 ref_p { C.to_world[3], C.to_world[7], C.to_world[11] };
        V3 d = uniform_sphere(s2);
        float radius = std::fmax(e.bs_radius, norm(ref_p - e.c)), dist = 2.f * radius, pdf = .25f * InvPi;
        float w = (pdf > 0.f ? rcp(pdf) : 0.f) * sqr(dist);
        emitter_weight = { w, w, w };
        v.p = fmadd3(d, dist, ref_p); v.n = -d; frame_from_normal(v, v.n); v.uv = s2; v.shape = -1;
    } else {
        float w; emitter_sample_position(e, s2, v.p, v.n, w);
        emitter_weight = { w, w, w };
        frame_from_normal(v, v.n); v.shape = e.shape;
    }
    V2 s3; s3.x = rng.next_float32(); s3.y = rng.next_float32();
    SensorDS sds = sensor_sample_direction(C, v.p, s3);
    v.wi = sds.d;
    (void) rng.next_float32();                                  // the wavelength sample (unused in RGB variants)
    // evaluated in the reference's order: emitter_idx_weight * emitter_weight * wav_weight * sensor_weight
    V3 weight = ((emitter_weight * idx_weight) * e.radiance) * sds.weight;
    connect_sensor(C, v, sds, nullptr, weight, block);
}

// ptracer.cpp:170-186 + scene.cpp:260-296 + area.cpp:92-115 / constant.cpp:93-111
static bool prepare_ray(const PtCtx &C, PCG32 &rng, Ray &ray, V3 &throughput) {
    const Scene &s = *C.s;
    if (C.shutter_open_time > 0.f) (void) rng.next_float32();
    float sample1 = rng.next_float32();
    V2 sample2, sample3;
    sample2.x = rng.next_float32(); sample2.y = rng.next_float32();
    sample3.x = rng.next_float32(); sample3.y = rng.next_float32();
    size_t count = s.emitters.size();
    if (count == 0) { throughput = { 0, 0, 0 }; return false; }
    int index = 0; float idx_weight = 1.f;
    if (count > 1) {
        float cf = (float) count, scaled = sample1 * cf;
        index = (int) std::min((uint32_t) scaled, (uint32_t) count - 1u); idx_weight = cf;
    }
    const Emitter &e = s.emitters[index];
    if (e.kind == 2) {
        V3 v0 = uniform_sphere(sample2);
        V3 orig = fmadd3(v0, e.bs_radius, e.c);
        V3 v1 = square_to_cosine_hemisphere(sample3);
        PtVertex f; frame_from_normal(f, -v0);
        ray.o = orig; ray.d = f.to_world(v1); ray.maxt = Largest;
        float area = 4.f * Pi * sqr(e.bs_radius);
        throughput = e.radiance * (area * Pi);
    } else if (e.directional) {
        // DirectionalArea::sample_ray (directionalarea.cpp:86-110): a point on the shape, direction = its normal, weight = area * radiance
        PtVertex v; float w; emitter_sample_position(e, sample2, v.p, v.n, w);
        throughput = e.radiance * w;
        ray = spawn_ray(v.p, v.n, v.n);
    } else {
        PtVertex v; float w; emitter_sample_position(e, sample2, v.p, v.n, w);
        frame_from_normal(v, v.n);
        V3 local = square_to_cosine_hemisphere(sample3);
        throughput = (e.radiance * w) * Pi;                       // pos_weight * wav_weight * Pi
        ray = spawn_ray(v.p, v.n, v.to_world(local));
    }
    if (count > 1) throughput = throughput * idx_weight;
    return true;
}

// BSDF::sample in TransportMode::Importance: as eval_pdf_sample's sample half, without the eta^2 factor of a refraction
// (dielectric.cpp:363-368)
static BsdfOut bsdf_sample_importance(const Scene &s, const Bsdf &b, V2 uv, V3 wi, float s1, V2 s2) {
    BsdfOut o = eval_pdf_sample(s, b, uv, wi, V3 { 0, 0, 1 }, s1, s2);
    if (b.type == 2 && o.seta != 1.f) {      // the transmitted lobe was selected (eta_it != 1)
        float r_i, ct, eit, eti; fresnel(wi.z, b.eta_s, r_i, ct, eit, eti);
        o.weight = b.trans;
    }
    return o;
}

// ptracer.cpp:205-283. `scalar`: the scalar variants draw a roulette number only once the roulette is on
static void trace_light_ray(const PtCtx &C, PCG32 &rng, Ray ray, V3 throughput, double *block) {
    const Scene &s = *C.s;
    float eta = 1.f; int depth = 1;
    bool active = (throughput.x != 0.f || throughput.y != 0.f || throughput.z != 0.f);
    if (!active) return;
    Hit h = intersect(s, ray, false);
    SI si = surface_interaction(s, ray, h);
    active = si.valid();
    if (C.max_depth >= 0) active = active && depth < C.max_depth;
    while (active) {
        const Bsdf &bsdf = s.bsdfs[s.shape_bsdf[si.shape]];
        V2 ss; ss.x = rng.next_float32(); ss.y = rng.next_float32();
        SensorDS sds = sensor_sample_direction(C, si.p, ss);
        PtVertex v; v.p = si.p; v.n = si.n; v.sh_s = si.sh_s; v.sh_t = si.sh_t; v.sh_n = si.sh_n; v.wi = si.wi; v.uv = si.uv; v.shape = si.shape;
        if (g_trace) std::printf("depth %d p %.9g %.9g %.9g shape %d wi %.6g %.6g %.6g thr %.6g %.6g %.6g | sds valid %d uv %.6g %.6g w %.6g pdf %.6g\n", depth, si.p.x, si.p.y,
                                 si.p.z, si.shape, si.wi.x, si.wi.y, si.wi.z, throughput.x, throughput.y, throughput.z, (int) sds.valid, sds.uv.x, sds.uv.y, sds.weight, sds.pdf);
        connect_sensor(C, v, sds, &bsdf, throughput * sds.weight, block);
        // bsdf->sample(ctx, si, sampler->next_1d(), sampler->next_2d()) (ptracer.cpp:243-245): the two draws are function
        // ARGUMENTS, whose evaluation order C++ leaves to the compiler; GCC (the pinned build, and any Linux build of the
        // reference) evaluates them right to left — the 2D sample is drawn first
        V2 s2; s2.x = rng.next_float32(); s2.y = rng.next_float32(); float s1 = rng.next_float32();
        BsdfOut bs = bsdf_sample_importance(s, bsdf, si.uv, si.wi, s1, s2);
        V3 wo_world = si.to_world(bs.wo);
        float wi_dot_geo_n = dot(si.n, -ray.d), wo_dot_geo_n = dot(si.n, wo_world);
        active = active && (wi_dot_geo_n * si.wi.z > 0.f) && (wo_dot_geo_n * bs.wo.z > 0.f);
        float correction = std::fabs((si.wi.z * wo_dot_geo_n) / (bs.wo.z * wi_dot_geo_n));
        if (g_trace) std::printf("   bs.wo9 %.9g %.9g %.9g wo_w %.9g %.9g %.9g s %.9g %.9g %.9g t %.9g %.9g %.9g n %.9g %.9g %.9g wi %.9g %.9g %.9g\n", bs.wo.x, bs.wo.y, bs.wo.z, wo_world.x, wo_world.y, wo_world.z,
                                 si.sh_s.x, si.sh_s.y, si.sh_s.z, si.sh_t.x, si.sh_t.y, si.sh_t.z, si.sh_n.x, si.sh_n.y, si.sh_n.z, si.wi.x, si.wi.y, si.wi.z);
        if (g_trace) std::printf("   bs.wo %.6g %.6g %.6g w %.6g %.6g %.6g corr %.6g active %d\n", bs.wo.x, bs.wo.y, bs.wo.z, bs.weight.x, bs.weight.y, bs.weight.z, correction, (int) active);
        throughput = throughput * (bs.weight * correction);
        eta *= bs.seta;
        active = active && (throughput.x != 0.f || throughput.y != 0.f || throughput.z != 0.f);
        if (!active) break;
        ray = spawn_ray(si.p, si.n, wo_world);
        h = intersect(s, ray, false);
        si = surface_interaction(s, ray, h);
        depth++;
        if (C.max_depth >= 0) active = active && depth < C.max_depth;
        active = active && si.valid();
        bool use_rr = depth > C.rr_depth;
        // sampler->next_1d(active): a masked draw does not advance the generator of an inactive lane (PCG32::next_uint32(mask),
        // ext/drjit/include/drjit/random.h). Scalar variants reach the draw only when the roulette applies (any_or<true>
        // evaluates the mask); JIT lanes reach it in every iteration
        if (use_rr || !C.scalar) {
            float q = std::fmin(std::fmax(throughput.x, std::fmax(throughput.y, throughput.z)) * sqr(eta), .95f);
            float r = active ? rng.next_float32() : 0.f;
            if (use_rr) { active = active && r < q; throughput = throughput * rcp(q); }
        }
    }
}

// ptracer.cpp:77-90
static void ptracer_sample(const PtCtx &C, PCG32 &rng, double *block) {
    if (C.max_depth != 0 && !C.hide_emitters) sample_visible_emitters(C, rng, block);
    Ray ray; V3 throughput;
    if (!prepare_ray(C, rng, ray, throughput)) return;
    trace_light_ray(C, rng, ray, throughput, block);
}

static void pt_setup(PtCtx &C, void *scene_h, void *cam_h, const float *to_world16, const float *to_camera16, int filter, float filter_param, int max_depth, int rr_depth,
                     int hide_emitters, float shutter_open_time, int cw, int ch, int cx, int cy) {
    C.s = (const Scene *) scene_h; C.cam = (const Camera *) cam_h;
    // The reference keeps an inverse next to every transform (callers that track it pass it in). For a camera placed with look_at (transform.h:254-282) it is
    // built analytically in float32: rows = the rotation's columns, translation = -(row . origin) accumulated as
    // ((r0 * -ox) + r1 * -oy) + r2 * -oz. Restated here for rigid matrices (film positions feed the 31-bin discretised
    // filter: an ulp in the camera-space point can move a splat weight by a bin); other matrices: double inverse, rounded once.
    const float *T = to_world16;
    for (int i = 0; i < 12; ++i) C.to_world[i] = T[i];
    float rows[3][3] = { { T[0], T[4], T[8] }, { T[1], T[5], T[9] }, { T[2], T[6], T[10] } };    // columns of the rotation
    bool rigid = true;
    for (int a = 0; a < 3 && rigid; ++a)
        for (int b = 0; b < 3; ++b) {
            double dp = (double) rows[a][0] * rows[b][0] + (double) rows[a][1] * rows[b][1] + (double) rows[a][2] * rows[b][2];
            if (std::fabs(dp - (a == b ? 1.0 : 0.0)) > 1e-5) rigid = false;
        }
    if (to_camera16) { for (int i = 0; i < 12; ++i) C.to_camera[i] = to_camera16[i]; }   // the transform's own float32 inverse
    else if (rigid) {
        float nx = -T[3], ny = -T[7], nz = -T[11];
        for (int r = 0; r < 3; ++r) {
            C.to_camera[4 * r] = rows[r][0]; C.to_camera[4 * r + 1] = rows[r][1]; C.to_camera[4 * r + 2] = rows[r][2];
            C.to_camera[4 * r + 3] = fmadd(rows[r][2], nz, fmadd(rows[r][1], ny, rows[r][0] * nx));
        }
    } else {
        double M[16], Mi[16];
        for (int i = 0; i < 16; ++i) M[i] = T[i];
        double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[6], g = M[8], hh = M[9], k = M[10];
        double det = a * (e * k - f * hh) - b * (d * k - f * g) + c * (d * hh - e * g);
        double R[9] = { (e * k - f * hh) / det, (c * hh - b * k) / det, (b * f - c * e) / det, (f * g - d * k) / det, (a * k - c * g) / det,
                        (c * d - a * f) / det, (d * hh - e * g) / det, (b * g - a * hh) / det, (a * e - b * d) / det };
        for (int r = 0; r < 3; ++r) {
            for (int q = 0; q < 3; ++q) Mi[4 * r + q] = R[3 * r + q];
            Mi[4 * r + 3] = -(R[3 * r] * M[3] + R[3 * r + 1] * M[7] + R[3 * r + 2] * M[11]);
        }
        for (int i = 0; i < 12; ++i) C.to_camera[i] = (float) Mi[i];
    }
    C.shutter_open_time = shutter_open_time; C.max_depth = max_depth; C.rr_depth = rr_depth; C.hide_emitters = hide_emitters != 0;
    C.flt.init(filter, filter_param); C.cw = cw; C.ch = ch; C.cx = cx; C.cy = cy;
}
// the perspective sensor's projection data (cam == nullptr): camera -> sample matrix of sensor.h:228-265 composed in double and
// rounded once, image rectangle and normalisation of perspective.cpp:186-194
static void pt_setup_persp(PtCtx &C, int w, int h) {
    if (C.cam) return;
    C.near_clip = g_pt_persp.near_clip; C.far_clip = g_pt_persp.far_clip;
    const double W = w, H = h, aspect = W / H, relx = C.cw / W, rely = C.ch / H, offx = C.cx / W, offy = C.cy / H;
    const double n_ = C.near_clip, f_ = C.far_clip, recip = 1.0 / (f_ - n_), cot = 1.0 / std::tan(0.5 * (double) g_pt_persp.fov * 3.14159265358979323846 / 180.0);
    auto mul = [](const double a[16], const double b[16], double out[16]) {
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { double v = 0; for (int k = 0; k < 4; ++k) v += a[4 * i + k] * b[4 * k + j]; out[4 * i + j] = v; }
    };
    const double P[16] = { cot, 0, 0, 0, 0, cot, 0, 0, 0, 0, f_ * recip, -n_ * f_ * recip, 0, 0, 1, 0 };
    const double T2[16] = { 1, 0, 0, -1, 0, 1, 0, -1.0 / aspect, 0, 0, 1, 0, 0, 0, 0, 1 };
    const double S2[16] = { -0.5, 0, 0, 0, 0, -0.5 * aspect, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
    const double T1[16] = { 1, 0, 0, -offx, 0, 1, 0, -offy, 0, 0, 1, 0, 0, 0, 0, 1 };
    const double S1[16] = { 1.0 / relx, 0, 0, 0, 0, 1.0 / rely, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
    double a[16], b[16], c[16], m[16];
    mul(T2, P, a); mul(S2, a, b); mul(T1, b, c); mul(S1, c, m);
    for (int i = 0; i < 16; ++i) C.c2s[i] = (float) m[i];
    // image rectangle at z = 1: sample_to_camera * (0,0,0) and (1,1,0), divided by their z (the near plane)
    Persp Pp; Pp.near_clip = C.near_clip; Pp.far_clip = C.far_clip; persp_setup(Pp, g_pt_persp.fov, w, h, C.cw, C.ch, C.cx, C.cy);
    const float x0 = Pp.s2c[3] / Pp.s2c[11], y0 = Pp.s2c[7] / Pp.s2c[11];
    const float x1 = (Pp.s2c[0] + Pp.s2c[3]) / Pp.s2c[11], y1 = (Pp.s2c[5] + Pp.s2c[7]) / Pp.s2c[11];
    C.rect_min[0] = std::min(x0, x1); C.rect_max[0] = std::max(x0, x1); C.rect_min[1] = std::min(y0, y1); C.rect_max[1] = std::max(y0, y1);
    C.normalization = 1.f / ((C.rect_max[0] - C.rect_min[0]) * (C.rect_max[1] - C.rect_min[1]));
    C.res[0] = (float) C.cw; C.res[1] = (float) C.ch;
}
}  // namespace orc

extern "C" {
// the perspective sensor (x fov in degrees, clip planes) used by the next particle-tracer calls that pass cam_h == NULL
void orc_ptracer_perspective(float fov, float near_clip, float far_clip) { orc::g_pt_persp.fov = fov; orc::g_pt_persp.near_clip = near_clip; orc::g_pt_persp.far_clip = far_clip; }
// PerspectiveCamera::sample_direction for n WORLD-space points: out11 = uv[2], p[3] (world), d[3], dist, pdf, weight
void orc_persp_sample_direction(float fov, float near_clip, float far_clip, const float *to_world16, int w, int h, int n, const float *p3, float *out11) {
    using namespace orc;
    g_pt_persp.fov = fov; g_pt_persp.near_clip = near_clip; g_pt_persp.far_clip = far_clip;
    PtCtx C; pt_setup(C, nullptr, nullptr, to_world16, nullptr, 1, 0.f, 4, 5, 0, 0.f, w, h, 0, 0);
    pt_setup_persp(C, w, h);
    for (int i = 0; i < n; ++i) {
        SensorDS ds = sensor_sample_direction(C, V3 { p3[3 * i], p3[3 * i + 1], p3[3 * i + 2] }, V2 { .5f, .5f });
        float *o = out11 + 11 * i;
        for (int k = 0; k < 11; ++k) o[k] = 0.f;
        if (!ds.valid) continue;
        o[0] = ds.uv.x; o[1] = ds.uv.y; o[2] = ds.p.x; o[3] = ds.p.y; o[4] = ds.p.z; o[5] = ds.d.x; o[6] = ds.d.y; o[7] = ds.d.z;
        o[8] = ds.dist; o[9] = ds.pdf; o[10] = ds.weight;
    }
}
// n_samples consecutive ParticleTracerIntegrator::sample calls drawn from ONE scalar sampler stream seeded with `seed`
// (sampler.cpp:131), sample_scale 1, into storage64 [ch, cw, 4] — the unit the reference exposes to Python as integrator.sample()
void orc_ptracer_sample(void *scene_h, void *cam_h, const float *to_world16, const float *to_camera16, int w, int h, int filter, float filter_param, uint32_t seed,
                        int n_samples, int max_depth, int rr_depth, int hide_emitters, float shutter_open_time, double *storage64);

// AdjointIntegrator::render. scalar != 0: the scalar branch (integrator.cpp:600-672) — ranges of grain_size =
// max(total / (4 n_threads), 1) samples, range k drawn from ONE stream seeded with  seed * (total / grain_size) + k. scalar == 0:
// the JIT branch (:673-740) — one stream per sample, seeded like the path integrator's wavefront lanes.
// storage64 [ch, cw, 4]: R, G, B and the (unused) weight channel; develop = the RGB channels as they are.
void orc_ptracer_render(void *scene_h, void *cam_h, const float *to_world16, const float *to_camera16, int w, int h, int cw, int ch, int cx, int cy, int filter,
                        float filter_param, uint32_t base_seed, uint32_t seed, int spp, int max_depth, int rr_depth, int hide_emitters,
                        float shutter_open_time, int scalar, int n_threads, double *storage64) {
    PtCtx C; pt_setup(C, scene_h, cam_h, to_world16, to_camera16, filter, filter_param, max_depth, rr_depth, hide_emitters, shutter_open_time, cw, ch, cx, cy);
    pt_setup_persp(C, w, h);
    C.sample_scale = (float) (cw * ch) / (float) ((uint32_t) spp * (uint32_t) (w * h));
    C.scalar = scalar != 0;
    size_t total = (size_t) w * h * spp, npix = (size_t) cw * ch;
    if (C.s->emitters.empty()) return;
    if (scalar) {
        size_t grain = std::max(total / (4 * (size_t) n_threads), (size_t) 1);
        uint32_t seed_eff = seed * ((uint32_t) total / (uint32_t) grain);
        size_t n_ranges = (total + grain - 1) / grain;
#pragma omp parallel
        {
            std::vector<double> local(npix * 4, 0.0);
#pragma omp for schedule(dynamic, 1)
            for (long long r = 0; r < (long long) n_ranges; ++r) {
                size_t begin = (size_t) r * grain, end = std::min(begin + grain, total);
                PCG32 rng; rng.seed((uint64_t) (base_seed + seed_eff + (uint32_t) begin / (uint32_t) grain), 0xda3e39cb94b95bdbULL);
                for (size_t i = begin; i < end; ++i) ptracer_sample(C, rng, local.data());
            }
#pragma omp critical
            for (size_t i = 0; i < npix * 4; ++i) storage64[i] += local[i];
        }
    } else {
#pragma omp parallel
        {
            std::vector<double> local(npix * 4, 0.0);
#pragma omp for schedule(dynamic, 4096)
            for (long long idx = 0; idx < (long long) total; ++idx) {
                PCG32 rng; sampler_seed_jit(rng, base_seed + seed, (uint32_t) idx);
                ptracer_sample(C, rng, local.data());
            }
#pragma omp critical
            for (size_t i = 0; i < npix * 4; ++i) storage64[i] += local[i];
        }
    }
}
void orc_ptracer_sample(void *scene_h, void *cam_h, const float *to_world16, const float *to_camera16, int w, int h, int filter, float filter_param, uint32_t seed,
                        int n_samples, int max_depth, int rr_depth, int hide_emitters, float shutter_open_time, double *storage64) {
    PtCtx C; pt_setup(C, scene_h, cam_h, to_world16, to_camera16, filter, filter_param, max_depth, rr_depth, hide_emitters, shutter_open_time, w, h, 0, 0);
    pt_setup_persp(C, w, h);
    C.sample_scale = 1.f; C.scalar = true;
    PCG32 rng; rng.seed((uint64_t) seed, 0xda3e39cb94b95bdbULL);
    for (int i = 0; i < n_samples; ++i) {
        PCG32 probe = rng;
        ptracer_sample(C, rng, storage64);
        if (g_trace) { int used = 0; while (!(probe.state == rng.state) && used < 100000) { (void) probe.next_float32(); ++used; } std::printf("sample %d: %d draws\n", i, used); }
    }
}
}
