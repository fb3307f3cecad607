/* lensrt.h — C ABI of liblensrt.so, the B200-native replacement for the render hot path of
 * gitmon/mitsuba3-camera (realistic-lens Sensor::sample_ray -> path integrator -> ImageBlock::put, and
 * its adjoint). Plain pointers and sizes only; no C++/torch types. File:line citations name the
 * reference interface each entry point stands in for (paths relative to the reference checkout).
 *
 * Conventions
 *  - every function returns 0 on success, a negative LRT_E* code otherwise; lrt_last_error() gives the
 *    message of the calling thread's last failure. Nothing throws across the boundary.
 *  - `*_dev` pointers are CUDA device pointers owned by the caller (e.g. torch tensors); `stream` is a
 *    cudaStream_t passed as void* (NULL = default stream). Calls are asynchronous on that stream unless
 *    the name ends in `_host` (host buffers in/out, copies and a stream sync inside the call).
 *  - handles are opaque, created/destroyed by the library, usable from one host thread at a time.
 *  - there is NO CPU fallback: every compute entry point fails with LRT_ENODEVICE without a CUDA device.
 */
#ifndef LENSRT_H
#define LENSRT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define LRT_OK 0
#define LRT_EINVAL (-1)
#define LRT_ENODEVICE (-2)
#define LRT_ECUDA (-3)
#define LRT_ELENS (-4) /* lens construction failed (paraxial focus ray blocked: realisticlens.cpp:2132-2154) */

typedef struct lrt_sensor lrt_sensor;
typedef struct lrt_scene lrt_scene;

const char *lrt_last_error(void);
int lrt_version(void);
/* number of kernels launched by this process so far (bench.py's gpu_launches) */
uint64_t lrt_launch_count(void);
int lrt_device_count(void);
int lrt_set_device(int device);
/* Measured ceilings of the current device for the roofline report (the path is neither HBM- nor tensor-bound; SURVEY.md 8d
 * asks for an FMA micro-kernel): out6 = FP32 TFLOP/s, G warp-instructions issued /s, L1-hit GB/s, L2-hit GB/s,
 * G shared-memory float atomics /s, SM clock MHz during the run. A few milliseconds per figure; synchronous. */
int lrt_measure_peaks(double *out6, void *stream);

/* ---------------------------------------------------------------------------------------- film geometry
 * Film::size/crop_size/crop_offset + the fork's physical size (include/mitsuba/render/film.h:193-198,
 * src/render/film.cpp:46-48: `diagonal` in millimetres). */
typedef struct lrt_film_geom {
    int32_t width, height;               /* full film resolution */
    int32_t crop_w, crop_h, crop_x, crop_y;
    float diagonal_mm;                   /* default 35 */
} lrt_film_geom;

/* ------------------------------------------------------------------------------------------- sensors
 * RealisticLensCamera constructor (src/sensors/realisticlens.cpp:832-933).
 * Prescription tables are front (world) -> rear (film), millimetres, as build_lens_from_data takes them
 * (:1568-1651): kinds (0 sphere, 1 plane, 2 aperture, 3 asphere), thickness, curvature radius,
 * semi-diameter, conic constant, 7 even-asphere coefficients per surface; `mats7` has n+1 rows
 * [model, p0..p5] (0: Cauchy A,B; 1: Sellmeier B1..3,C1..3), world side first. `scale` multiplies all
 * lengths (exp2*: 35/1.57), `stop_ratio` the aperture semi-diameter (:850,:1596). `fine_focus` runs the
 * thick-lens autofocus for `object_distance` metres (:902-908, :2107-2171).
 * to_world16: row-major 4x4 camera-to-world (no scale allowed, :833). */
int lrt_sensor_create_prescription(int32_t n, const int32_t *kinds, const float *thickness_mm,
                                   const float *radius_mm, const float *semi_diameter_mm, const float *kappa,
                                   const float *ai7, const float *mats7, float scale, float stop_ratio,
                                   int32_t fine_focus, float object_distance, const lrt_film_geom *film,
                                   const float *to_world16, float near_clip, float far_clip,
                                   lrt_sensor **out);
/* The same camera as a SPECTRAL variant builds it (the fork's `*_spectral*` dispersion mode): identical, except that the
 * autofocus traces rays of Wavelength(589.3f) nm (:2122-2146), so the nominal IORs of the table are taken at 0.001f * 589.3f um —
 * one ulp above the rgb variants' 0.5893f. Use with the wavelength-aware entry points below. */
int lrt_sensor_create_prescription_spectral(int32_t n, const int32_t *kinds, const float *thickness_mm,
                                            const float *radius_mm, const float *semi_diameter_mm, const float *kappa,
                                            const float *ai7, const float *mats7, float scale, float stop_ratio,
                                            int32_t fine_focus, float object_distance, const lrt_film_geom *film,
                                            const float *to_world16, float near_clip, float far_clip,
                                            lrt_sensor **out);
/* Same camera from ready-made interfaces (rear -> front, metres), rows of 16 floats:
 * [kind, z, semi_diameter, R (sphere) | c,K,r_scale (asphere, unitless as :557-572), -, a0..a6 at [6..12],
 *  ior_left (film side), ior_right, -]. Used for the parametric designs (:936-975, :1322-1370) and for
 * optimisation loops that own the lens parameters. */
int lrt_sensor_create_interfaces(int32_t n, const float *rows16, float film_z, int32_t fine_focus,
                                 float object_distance, const lrt_film_geom *film, const float *to_world16,
                                 float near_clip, float far_clip, lrt_sensor **out);
/* PerspectiveCamera (src/sensors/perspective.cpp): x-axis field of view in degrees. */
int lrt_sensor_create_perspective(float x_fov_deg, const lrt_film_geom *film, const float *to_world16,
                                  float near_clip, float far_clip, lrt_sensor **out);
void lrt_sensor_destroy(lrt_sensor *s);
int lrt_sensor_num_interfaces(const lrt_sensor *s);
/* current interface table (after autofocus), rows16 as above */
int lrt_sensor_get_interfaces(const lrt_sensor *s, float *rows16);
/* parameters_changed(): replace the interface table / pose without re-running the autofocus */
int lrt_sensor_set_interfaces(lrt_sensor *s, int32_t n, const float *rows16);
/* the same with the table in DEVICE memory (rows16_dev [n,16]): one small kernel on `stream`, no host round trip — for design
 * loops whose variables live on the GPU. The host copy is refreshed lazily by the getters / *_host entry points. */
int lrt_sensor_set_interfaces_dev(lrt_sensor *s, int32_t n, const float *rows16_dev, void *stream);
int lrt_sensor_set_to_world(lrt_sensor *s, const float *to_world16);
int lrt_sensor_get_info(const lrt_sensor *s, float *sample_to_film16, float *focus_delta, float *film_z);

/* Exit-pupil sampling (the sensor's `sample_exit_pupil` property, src/sensors/realisticlens.cpp:913-921).
 * enable != 0 builds the 64-segment LUT on the GPU -- compute_exit_pupil_bounds / bound_exit_pupil (:1685-1809):
 * 64 x 2^20 Halton rays (RadicalInverse::eval, include/mitsuba/core/qmc.h:68-91) traced film -> world -- and switches
 * sample_ray to sample_exit_pupil (:1812-1845) with the weight shutter_open_time * box area / rear_z^2 (:2279-2289).
 * shutter_open_time is Sensor::shutter_open_time() (0 unless the scene sets shutter_open/close, as in the reference).
 * set_interfaces() rebuilds the LUT while the mode is on. get: bounds256 = 64 x (min.x, min.y, max.x, max.y), metres. */
int lrt_sensor_set_exit_pupil(lrt_sensor *s, int32_t enable, float shutter_open_time, void *stream);
int lrt_sensor_get_exit_pupil(const lrt_sensor *s, float *bounds256);
/* RadicalInverse::eval(base_index, index) for base_index 0..2 (primes 2, 3, 5); host function, -1 if out of range */
float lrt_radical_inverse(int32_t base_index, uint64_t index);

/* Sensor::sample_ray (realisticlens.cpp:2209-2292) for n samples: pos2/ap2 = film and aperture samples in
 * [0,1]^2; out10 = origin[3], direction[3], maxt, weight[3] (world space). */
int lrt_sensor_sample_ray(const lrt_sensor *s, int64_t n, const float *pos2_dev, const float *ap2_dev,
                          float *out10_dev, void *stream);
int lrt_sensor_sample_ray_host(const lrt_sensor *s, int64_t n, const float *pos2, const float *ap2, float *out10);
/* Dispersion (the spectral variants of the fork: DispersiveMaterial::compute_ior(ray), realisticlens.cpp:169-183,226-240):
 * every ray carries a wavelength in nm (the FIRST entry of the reference's wavelength vector) and each interface refracts
 * with the IORs of its two media at that wavelength — Cauchy A + B / w^2 or Sellmeier sqrt(1 + sum B_i w^2 / (w^2 - C_i)),
 * w = 0.001 * wavelength in um. Sensors built from a prescription carry their media; sensors built from ready-made
 * interfaces need lrt_sensor_set_materials: per interface (rear -> front) two rows of 7 floats, left (film side) then right,
 * [0, A, B, -, -, -, -] or [1, B1, B2, B3, C1, C2, C3].
 * lrt_sensor_sample_ray_wl = Sensor::sample_ray of a spectral variant, geometric part: out10 as lrt_sensor_sample_ray (the
 * caller multiplies the weight by sample_wavelengths()' own). lrt_sensor_trace_wl = trace_ray_from_film / _from_world with the
 * ray's wavelength, on the device: od6 / out7 as lrt_sensor_trace_host. */
/* compute_exit_pupil_bounds as a SPECTRAL variant runs it (realisticlens.cpp:1718-1727): every ray of the LUT build carries the
 * wavelength lerp(380, 700, RadicalInverse(2, i)) nm. Needed before the _wl entry points may sample the exit pupil. */
int lrt_sensor_set_exit_pupil_spectral(lrt_sensor *s, int32_t enable, float shutter_open_time, void *stream);
int lrt_sensor_set_materials(lrt_sensor *s, int32_t n, const float *mat_lr14);
int lrt_sensor_get_materials(const lrt_sensor *s, float *mat_lr14);
int lrt_sensor_sample_ray_wl(const lrt_sensor *s, int64_t n, const float *pos2_dev, const float *ap2_dev,
                             const float *wavelength_nm_dev, float *out10_dev, void *stream);
int lrt_sensor_trace_wl(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6_dev,
                        const float *wavelength_nm_dev, float *out7_dev, void *stream);
int lrt_sensor_trace_wl_host(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6,
                             const float *wavelength_nm, float *out7);
/* trace_ray_from_film / trace_ray_from_world (:1654-1683 / :1960-1984) in camera space:
 * od6 = o[3], d[3]; out7 = o[3], d[3], active (1/0). */
int lrt_sensor_trace_host(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6, float *out7);
/* Sensor::sample_direction (realisticlens.cpp:2301-2394) for n reference points given in CAMERA space (apply
 * to_world^-1 first); s2 = aperture samples; out11 = uv[2] (pixels of the crop window), p[3], d[3], dist, pdf,
 * weight (camera space). Invalid connections return zeros. */
int lrt_sensor_sample_direction(const lrt_sensor *s, int64_t n, const float *p3_dev, const float *s2_dev,
                                float *out11_dev, void *stream);
int lrt_sensor_sample_direction_host(const lrt_sensor *s, int64_t n, const float *p3, const float *s2, float *out11);
/* Host vector-Jacobian product of one lens trace (chief-ray sensitivities; the reference has no counterpart:
 * its lens tables are constants, its notebooks differentiate mesh lenses through Dr.Jit,
 * notebooks/scripts/lens_design/lens.py:67-85). od6 as lrt_sensor_trace_host, bar6 = adjoint of the outgoing
 * ray; out_bar6 = adjoint of the incoming ray; d/d(rows16) is ADDED to grad_rows16 [n_if,16] (same column
 * layout as rows16: z at 1, R | c at 3, K at 4, a0..a6 at 6..12, ior_left 13, ior_right 14). */
int lrt_sensor_trace_vjp_host(const lrt_sensor *s, int32_t from_world, int64_t n, const float *od6, const float *bar6,
                              float *out_bar6, float *grad_rows16, int32_t *active);

/* --------------------------------------------------------------------------------------------- sampler
 * IndependentSampler over per-sample PCG32 streams (src/samplers/independent.cpp:88-97,
 * src/render/sampler.cpp:115-134 JIT branch): out[i*ndraw + k] = k-th next_1d() of wavefront lane idx0+i. */
int lrt_sampler_draw(uint32_t seed, uint32_t idx0, int64_t n, int32_t ndraw, float *out_dev, void *stream);
int lrt_sampler_draw_host(uint32_t seed, uint32_t idx0, int64_t n, int32_t ndraw, float *out);

/* ------------------------------------------------------------------------------------------------ film
 * ImageBlock (include/mitsuba/render/imageblock.h:52-72 ctor arguments) + reconstruction filter. */
typedef struct lrt_film_desc {
    int32_t size_x, size_y;   /* block size = film crop size */
    int32_t off_x, off_y;     /* block offset = film crop offset */
    int32_t border;           /* border pixels on each side (0 when `border=false`) */
    int32_t channels;         /* 4 for hdrfilm rgb: R,G,B,W (hdrfilm.cpp:245-273) */
    int32_t filter;           /* 0 box, 1 tent, 2 gaussian */
    float filter_param;       /* tent radius / gaussian stddev; 0 = plugin default (1 / 0.5) */
    int32_t coalesce;         /* imageblock.cpp:418-531 window (render: spp >= 4, integrator.cpp:290) */
    int32_t normalize;        /* imageblock.cpp:318-342 / :461-474 (ptracer splats) */
} lrt_film_desc;

/* ImageBlock::put (src/render/imageblock.cpp:174-532) for n samples: pos2 in pixel coordinates of the full
 * film, values[n*channels]; accumulates (atomically) into accum [(size_y+2b),(size_x+2b),channels] f32. */
int lrt_film_put(const lrt_film_desc *film, int64_t n, const float *pos2_dev, const float *values_dev,
                 float *accum_dev, void *stream);
/* HDRFilm::develop for pixel_format=rgb (src/films/hdrfilm.cpp:304-395): storage [npix,4] -> rgb [npix,3]. */
int lrt_film_develop_rgb(int64_t npix, const float *storage4_dev, float *rgb_dev, void *stream);
/* ReconstructionFilter::eval (host; same code as the kernels) */
int lrt_filter_eval_host(int32_t filter, float param, int32_t n, const float *x, float *out, float *radius);

/* ------------------------------------------------------------------------------------------------ scene
 * Flat scene description (what mi.load_dict assembles from `rectangle` / `cube` / mesh shapes, `diffuse` /
 * `conductor` / `dielectric` (+`twosided`) BSDFs, `bitmap` textures and `area` emitters). All geometry is triangles;
 * area emitters sit on rectangles (centre + two half-edge vectors, src/shapes/rectangle.cpp:115-185).
 *   bsdfs   [n_bsdfs,16]   : type (0 diffuse, 1 conductor, 2 dielectric), twosided, reflectance rgb (diffuse)
 *                            or specular_reflectance rgb, eta rgb, k rgb (conductor), specular_transmittance
 *                            rgb, eta = int_ior/ext_ior (dielectric), texture id + 1 of a diffuse `reflectance`
 *                            bitmap (0 = the constant rgb)
 *   emitters[n_emitters,16]: radiance rgb, shape id, centre xyz, half-edge u xyz, half-edge v xyz, kind (col 13), pad.
 *                            kind 0 = `area` on a `rectangle` (sampled through the rectangle's parameterisation, rectangle.cpp:166-185);
 *                            kind 1 = `area` on any mesh: faces sampled by area, Mesh::sample_position (src/render/mesh.cpp:430-470,
 *                            745-803); centre / half-edges are ignored, the shape's triangles (and vertex normals) are used;
 *                            kind 2 = `constant` environment emitter (src/emitters/constant.cpp): shape id -1, columns 4..6 the
 *                            centre and column 7 the (padded) radius of the scene's bounding sphere (constant.cpp:75-85). The
 *                            order of the rows is the order of Scene::emitters(): a sample picks row floor(u * n_emitters).
 * Textures: kind 0 = BitmapTexture (src/textures/bitmap.cpp:115-310), RGB float texels (already linear), bilinear filter;
 * kinds 1..4 = the procedural patterns `checkerboard` (src/textures/checkerboard.cpp:70-89) and the fork's `checkerlines`,
 * `comb2d`, `polarlines` (src/textures/{checkerlines,comb2d,polarlines}.cpp) with constant colours. */
typedef struct lrt_texture_desc {
    int32_t width, height;
    int32_t wrap;              /* wrap_mode: 0 repeat, 1 mirror, 2 clamp */
    const float *rgb;          /* [height, width, 3] host memory (kind 0) */
    float to_uv[6];            /* first two rows of the 3x3 `to_uv` transform (identity: 1 0 0 / 0 1 0) */
    int32_t kind;              /* 0 bitmap, 1 checkerboard, 2 checkerlines, 3 comb2d, 4 polarlines */
    float color0[3], color1[3];
    float linewidth;           /* kinds 2..4 */
    float spokes;              /* kind 4 */
} lrt_texture_desc;

typedef struct lrt_scene_desc {
    int32_t n_vertices; const float *vertices;       /* [n_vertices,3] world space */
    int64_t n_triangles; const int32_t *indices;     /* [n_triangles,3] */
    const int32_t *tri_shape;                        /* [n_triangles] shape id */
    int32_t n_shapes; const int32_t *shape_bsdf;     /* [n_shapes] */
    const int32_t *shape_emitter;                    /* [n_shapes], -1 = not an emitter */
    int32_t n_bsdfs; const float *bsdfs;
    int32_t n_emitters; const float *emitters;
    /* Mesh vertex attributes (src/render/mesh.cpp:1448-1500), all optional (NULL): shape_flags[s] bit 0 = the vertices of
     * shape s carry normals (interpolated shading normal), bit 1 = texture coordinates (si.uv and dp_du), bit 2 = the shape
     * is a `rectangle` given as the two triangles (0,1,2),(0,2,3) of its corners (-1,-1),(1,-1),(1,1),(-1,1): its shading
     * frame follows the rectangle's parameterisation (src/shapes/rectangle.cpp:115-123) with or without texture coordinates. */
    const float *vertex_normals;                     /* [n_vertices,3] */
    const float *vertex_uvs;                         /* [n_vertices,2] */
    const int32_t *shape_flags;                      /* [n_shapes] */
    int32_t n_textures; const lrt_texture_desc *textures;
    /* Analytic `sphere` shapes (src/shapes/sphere.cpp), intersected in double precision like the reference's scalar variants
     * and kept outside the BVH (tested one by one: meant for the handful of spheres scene files use). 32 floats each:
     * centre xyz, radius, shape id, flip_normals, 0, 0, then the three rows of to_world (12) and of to_object (12): the
     * spherical parameterisation (si.uv, dp_du) lives in object space. Such shapes own no triangles.
     * A `disk` (src/shapes/disk.cpp: the unit disk of object space's xy plane, float32 like the reference) rides in the same
     * table: floats [0..2] = to_world * Normal3f(0, 0, 1) (normalised by the library, disk.cpp:122), [3] = 0, [4] = shape id,
     * [6] = 1, then the same two matrices. An emitter on it is an emitter row of kind 3 ([13] = 3: [4..6] = the translation,
     * [7..9] and [10..12] = the first two columns of to_world), an `area` light or a `directionalarea` ([14] = 1). */
    int32_t n_spheres; const float *spheres;
} lrt_scene_desc;

/* Scene::Scene (src/render/scene.cpp:22-96): builds the acceleration structure (8-wide quantised BVH, on the
 * host) and uploads everything to the current device. */
int lrt_scene_create(const lrt_scene_desc *desc, lrt_scene **out);
/* SceneParameters.update() (src/python/python/util.py:292-353) for the differentiable parameters: replace the BSDF rows
 * ([n_bsdfs,16], NULL = keep) and / or the emitters' radiance ([n_emitters,3], NULL = keep). Geometry is untouched. */
int lrt_scene_set_params(lrt_scene *s, const float *bsdfs, const float *emitter_radiance3);
void lrt_scene_destroy(lrt_scene *s);
int lrt_scene_info(const lrt_scene *s, int64_t *n_nodes, int64_t *n_tris, int32_t *bvh_depth, double *build_seconds);
/* Mesh-lens mode (notebooks/scripts/lens_design/lens.py:343-565): the vertices of an existing scene moved (same topology).
 * vertices_dev [n_vertices,3] / normals_dev [n_vertices,3] (NULL: keep) are DEVICE pointers; the triangle records are rewritten
 * and the 8-wide BVH is refitted bottom-up on `stream` — the stand-in for Mesh::parameters_changed + the accelerator rebuild
 * (src/render/mesh.cpp:85-162). */
int lrt_scene_refit(lrt_scene *s, const float *vertices_dev, const float *normals_dev, void *stream);
/* Rays through refractive meshes: od6 = o[3], d[3]; shape_eta_dev [n_shapes] = int_ior / ext_ior of the interface a shape
 * models (interior = behind its normal; 0 = absorber, e.g. a baffle); at most max_interactions surfaces. out8 = o[3], d[3] of the
 * ray leaving the last surface, alive (1/0), product of the Fresnel transmittances: the transmission lobe of the `dispersive` /
 * `dielectric` BSDF (src/bsdfs/dispersive_dielectric.cpp:250-431) applied about the interpolated vertex normal. */
int lrt_scene_trace_refractive(const lrt_scene *s, int64_t n, const float *od6_dev, const float *shape_eta_dev,
                               int32_t max_interactions, float *out8_dev, void *stream);
/* The adjoint of lrt_scene_trace_refractive — what Dr.Jit's reverse mode yields for the notebooks' mesh lenses: Möller–Trumbore
 * differentiable in the vertex positions (include/mitsuba/render/mesh.h:432-452; Mesh::compute_surface_interaction,
 * src/render/mesh.cpp:1370-1500), the interpolated vertex normal, fresnel / refract (include/mitsuba/render/fresnel.h:35-84,311-315).
 * grad_out8 = d(loss)/d(out8) (o[3], d[3], -, weight). Writes grad_od6 [n,6] = d(loss)/d(od6); ACCUMULATES (float atomics; the
 * caller zeroes) grad_vertices [n_vertices,3], grad_normals [n_vertices,3] (the interpolated vertex normals as independent
 * inputs; their dependence on the positions, mesh.cpp:321-375, is chained by the caller) and grad_eta [n_shapes]; each of the
 * three may be NULL. Which triangle is hit, absorption and total internal reflection are detached, as in the reference's AD;
 * dead rays contribute nothing. max_interactions <= 32. */
int lrt_scene_trace_refractive_backward(const lrt_scene *s, int64_t n, const float *od6_dev, const float *shape_eta_dev,
                                        int32_t max_interactions, const float *grad_out8_dev, float *grad_od6_dev,
                                        float *grad_vertices_dev, float *grad_normals_dev, float *grad_eta_dev, void *stream);
/* Scene::ray_intersect_preliminary / ray_test (src/render/scene.cpp:181-210) for n rays:
 * od7 = o[3], d[3], maxt; out4 = t, u, v, primitive index (-1: miss; t = inf). */
int lrt_scene_intersect(const lrt_scene *s, int32_t any_hit, int64_t n, const float *od7_dev, float *out4_dev, void *stream);

/* ----------------------------------------------------------------------------------------------- render
 * Integrator::render (src/render/integrator.cpp:120-367, JIT sample order) with the `path` integrator
 * (src/integrators/path.cpp:95-287; max_depth, rr_depth as in src/render/integrator.cpp:765-790). */
typedef struct lrt_render_params {
    uint32_t seed;
    int32_t spp;
    int32_t max_depth;        /* -1 = unbounded */
    int32_t rr_depth;         /* default 5 */
    int32_t shard_index, shard_count;   /* pixel tiles are dealt round-robin over shards (multi-GPU) */
    int32_t chunks;           /* 0 = automatic split of a pixel's samples into work items */
    /* Sampler convention (src/render/sampler.cpp:115-134). 0: the JIT variants' — one PCG32 stream per sample,
     * (v0, v1) = tea(seed, idx) — the fast path. 1: the scalar variants' — one stream per pixel seeded with
     * base_seed + seed * W * H + block_id * block_size^2 + morton index (src/render/integrator.cpp:195-200,388-402), spiral
     * block order, discretised filter, scalar control flow of the path loop: a single-GPU parity mode whose images
     * match the reference's scalar_rgb variant sample for sample. */
    int32_t rng_mode;
    int32_t block_size;       /* rng_mode 1: render block size, 0 = 32 (MI_BLOCK_SIZE) */
    uint32_t base_seed;       /* rng_mode 1: the sampler's `seed` property (rng_mode 0 callers fold it into `seed`) */
    int32_t hide_emitters;    /* path.cpp:113-114: with an environment emitter, camera rays that escape yield no valid sample */
} lrt_render_params;

/* Accumulates R,G,B,W into accum_dev [size_y+2b, size_x+2b, 4] (caller zeroes it; ImageBlock storage of
 * hdrfilm; 16-byte aligned: pixels are flushed with 128-bit reductions). Asynchronous on `stream`. */
int lrt_render(const lrt_scene *scene, const lrt_sensor *sensor, const lrt_film_desc *film,
               const lrt_render_params *params, float *accum_dev, void *stream);
/* Adjoint of lrt_render for the scene parameters the path is linear in — the part of PRBIntegrator.render_backward
 * (src/python/python/ad/integrators/prb.py:59-257, common.py:631-796) that reaches diffuse `reflectance`, conductor /
 * dielectric `specular_reflectance`, dielectric `specular_transmittance` and area-light `radiance`. grad_storage_dev =
 * d(loss)/d(accum) [size_y+2b, size_x+2b, 4] (the RGB channels are used; W does not depend on parameters). ADDS to
 * grad_bsdfs_dev [n_bsdfs,16] (columns 2..4 and, for the transmittance, 11..13) and grad_emitters_dev [n_emitters,16]
 * (columns 0..2), laid out like the tables of lrt_scene_desc; the caller zeroes them. Same samples, streams and
 * Russian-roulette decisions as lrt_render with equal params (sampling and the roulette weight are detached, as in prb).
 * One pass: instead of prb's second replay pass, each path keeps (BSDF id, lobe, radiance gathered so far) per bounce —
 * at most 16 bounces: max_depth outside [0, 17] is an error (LRT_EINVAL), never a silent truncation. Textured reflectances
 * receive no gradient. */
int lrt_render_backward(const lrt_scene *scene, const lrt_sensor *sensor, const lrt_film_desc *film,
                        const lrt_render_params *params, const float *grad_storage_dev, float *grad_bsdfs_dev,
                        float *grad_emitters_dev, void *stream);
/* mi.render(scene) for host callers: render + HDRFilm::develop; rgb_out [size_y,size_x,3] host memory. */
int lrt_render_host(const lrt_scene *scene, const lrt_sensor *sensor, const lrt_film_desc *film,
                    const lrt_render_params *params, float *rgb_out);
/* ------------------------------------------------------------------------------------------- particle tracer
 * ParticleTracerIntegrator (src/integrators/ptracer.cpp:77-371) under AdjointIntegrator::render (src/render/integrator.cpp:
 * 544-760): every sample connects a point on an emitter to the sensor (sample_visible_emitters), then shoots a light path from
 * an emitter (Scene::sample_emitter_ray) and connects each of its vertices to the sensor through Sensor::sample_direction
 * (realistic lens: src/sensors/realisticlens.cpp:2301-2394) — the integrator the fork's lens-design notebooks render with. */
typedef struct lrt_ptrace_params {
    uint32_t seed;
    int32_t spp;              /* samples = film width * height * spp (the FULL film size, integrator.cpp:573-574) */
    int32_t max_depth;        /* -1 = unbounded */
    int32_t rr_depth;         /* default 5 */
    int32_t hide_emitters;    /* skip the directly visible emitters (ptracer.cpp:80-81) */
    int32_t shard_index, shard_count;   /* rng_mode 0: sample idx = shard_index + k * shard_count (multi-GPU) */
    /* 0: one PCG32 stream per sample, (v0, v1) = tea(seed, idx) — the JIT variants' convention, the fast path.
     * 1: the scalar variants' convention (integrator.cpp:600-640): ranges of max(total / (4 n_threads), 1) samples, range k drawn
     *    from ONE stream seeded with base_seed + seed * (total / grain) + k, discretised filter: images comparable sample for
     *    sample with a scalar_rgb build of the reference running n_threads threads. Single GPU, one thread per range. */
    int32_t rng_mode;
    int32_t n_threads;        /* rng_mode 1 */
    uint32_t base_seed;       /* the sampler's `seed` property */
    float to_camera[12];      /* rows of the inverse of the sensor's to_world as the scene description carries it (Transform4f keeps
                                 an inverse next to every matrix); all zero = invert to_world numerically */
} lrt_ptrace_params;
/* Accumulates R,G,B into accum_dev [size_y, size_x, 4] (caller zeroes it; the weight channel stays 0: the image is the RGB
 * channels as they are). film: channels 4, normalize 1, coalesce 0 (integrator.cpp:706-717), border 0. Asynchronous on `stream`.
 * lrt_render_stats afterwards: [0] light paths, [1] connections splatted, [2] rays traced. */
int lrt_ptrace(const lrt_scene *scene, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_ptrace_params *params,
               float *accum_dev, void *stream);

/* Lens adjoint of lrt_ptrace — what `dr.backward(loss)` through a `ptracer` render delivers to the lens parameters in the
 * notebooks' design loop (notebooks/scripts/lens_design/design_problem.py:495-584; generic Integrator::render_backward,
 * src/render/integrator.cpp:66-84), here for the analytic interface table. grad_accum_dev = d(loss)/d(accum) [size_y, size_x, 4]
 * (RGB used); ADDS d(loss)/d(rows16) to grad_rows16_dev [n_if, 16] (caller zeroes it; columns as in lrt_spot_render_backward).
 * The same light paths are replayed (rng_mode 0, equal params): every connection contributes the adjoint of its splat position and
 * of the sensor's importance weight; light-path vertices, BSDF values and visibility do not depend on the lens; vignetting masks
 * are piecewise constant and detached, as in the spot path. */
int lrt_ptrace_backward(const lrt_scene *scene, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_ptrace_params *params,
                        const float *grad_accum_dev, float *grad_rows16_dev, void *stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid allocation). lrt_render_host without any allocation:
 * render (rng_mode 0) + HDRFilm::develop with the R,G,B,W accumulator in a fixed 4 MiB device array inside the library, so the
 * film may have at most lrt_render_rgb_noalloc_max_pixels() = 2^18 pixels (e.g. 512 x 512), border 0, 4 channels; larger films
 * are refused with LRT_EINVAL. rgb_out_dev [size_y, size_x, 3] is DEVICE memory owned by the caller; asynchronous on `stream`.
 * Calls no cudaMalloc / cudaFree. The sensor must have been used once before (its first use allocates its device table), else
 * LRT_EINVAL. All calls share the one array (cleared at the start of each call): calls that overlap on the GPU race. */
int lrt_render_rgb_noalloc(const lrt_scene *scene, const lrt_sensor *sensor, const lrt_film_desc *film,
                           const lrt_render_params *params, float *rgb_out_dev, void *stream);
int64_t lrt_render_rgb_noalloc_max_pixels(void);
/* counters of the last render on this scene: samples, lens misses, rays traced, path-loop iterations,
 * BVH nodes fetched (80 B each), triangles tested (48 B each), traversal-stack clamp events (must be 0: a clamped stack
 * may miss the closest hit) */
int lrt_render_stats(const lrt_scene *scene, uint64_t *out7);

/* ------------------------------------------------------------------------------- lens-design spot path
 * Light-traced spot image of a point emitter through the realistic lens and its adjoint: the B200 form of
 * one `ptracer` render + render_backward of the notebooks' design loop
 * (notebooks/scripts/lens_design/design_problem.py:495-584; src/integrators/ptracer.cpp connect_sensor ->
 * Sensor::sample_direction realisticlens.cpp:2301-2394 -> ImageBlock::put normalize=true, coalesce=false,
 * src/render/integrator.cpp:706-717; generic Integrator::render_backward src/render/integrator.cpp:66-84). */
typedef struct lrt_spot_params {
    uint32_t seed;
    int32_t spp;                       /* size_x*size_y*spp samples in total; sample scale 1/spp (integrator.cpp:671-676) */
    float src_p[3];                    /* point emitter, CAMERA space (a distant point = a field angle) */
    float intensity[3];                /* RGB radiant intensity */
    int32_t shard_index, shard_count;  /* sample idx = shard_index + k*shard_count: identical streams for any GPU count */
    float origin_offset;               /* 0: rays start at src_p as in sample_direction (:2336). > 0: the same ray starts this
                                          many metres in front of the sampled front-plane point (well-conditioned in float32
                                          for distant sources; see lrt_lens.cuh) */
    int32_t fast;                      /* 0: strict float32 (every operation rounded as the reference's scalar build does: the
                                          miss masks are bit-exact). 1: the same algorithm with fused multiply-adds and hardware
                                          reciprocal / square-root approximations (last-ulp differences, ~2x the rays per second):
                                          what design loops use; primal and adjoint must use the same flavour */
    /* Optional device-side seed for loops captured in a CUDA graph (launch parameters are frozen at capture): when
     * seed_dev != NULL the effective seed is (seed + *seed_dev * seed_stride) ^ seed_xor, read by the kernel at run time. */
    const uint32_t *seed_dev;
    uint32_t seed_stride, seed_xor;
} lrt_spot_params;

/* Accumulates intensity*weight/spp (R,G,B) and 1/spp (W) times the normalised filter weights into accum_dev
 * [size_y+2b, size_x+2b, 4]; the spot image is channels 0..2 (AdjointIntegrator films are not weight-divided).
 * stats_dev (optional, 2 x u64, caller-zeroed): samples traced, samples that reached the film. */
int lrt_spot_render(const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_spot_params *params,
                    float *accum_dev, uint64_t *stats_dev, void *stream);
/* Adjoint: grad_accum_dev = d(loss)/d(accum) (same shape); ADDS d(loss)/d(rows16) to grad_rows16_dev [n_if,16]
 * (caller zeroes it). Samples, streams and vignetting masks are those of lrt_spot_render with equal params. */
int lrt_spot_render_backward(const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_spot_params *params,
                             const float *grad_accum_dev, float *grad_rows16_dev, uint64_t *stats_dev, void *stream);

#ifdef __cplusplus
}
#endif
#endif
