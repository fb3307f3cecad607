// lensrt — scene construction (host BVH build + upload) and the render entry points of include/lensrt.h.
#include "lrt_scene.h"
#include <chrono>
#include <cstring>
#include <cmath>

namespace lrt {
struct RenderArgs;
int launch_render(const lrt_scene *sc, lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                  float *accum_dev, const float *grad_storage, float *grad_bsdfs, float *grad_emitters, cudaStream_t stream);
int launch_render_scalar(const lrt_scene *sc, lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                         float *accum_dev, cudaStream_t stream);
}
using namespace lrt;

template <typename T> static int upload(void **dst, const T *src, size_t n) {
    size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
    LRT_CUDA(cudaMalloc(dst, bytes));
    if (n) LRT_CUDA(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
    return LRT_OK;
}

extern "C" {

int lrt_scene_create(const lrt_scene_desc *d, lrt_scene **out) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(d && out, "null argument");
    LRT_REQUIRE(d->n_triangles >= 0 && d->n_shapes >= 0 && d->n_bsdfs >= 0 && d->n_emitters >= 0, "negative count");
    LRT_REQUIRE(d->n_triangles == 0 || (d->vertices && d->indices && d->tri_shape), "null geometry");
    LRT_REQUIRE(d->n_triangles < (1 << 28), "too many triangles (2^28 max)");
    for (int64_t t = 0; t < d->n_triangles; ++t) {
        LRT_REQUIRE(d->tri_shape[t] >= 0 && d->tri_shape[t] < d->n_shapes, "tri_shape out of range");
        for (int k = 0; k < 3; ++k) LRT_REQUIRE(d->indices[3 * t + k] >= 0 && d->indices[3 * t + k] < d->n_vertices, "vertex index out of range");
    }
    for (int s = 0; s < d->n_shapes; ++s) {
        LRT_REQUIRE(d->shape_bsdf[s] >= 0 && d->shape_bsdf[s] < d->n_bsdfs, "shape_bsdf out of range");
        LRT_REQUIRE(d->shape_emitter[s] >= -1 && d->shape_emitter[s] < d->n_emitters, "shape_emitter out of range");
    }
    LRT_REQUIRE(d->n_textures >= 0 && (d->n_textures == 0 || d->textures), "bad texture table");
    for (int t = 0; t < d->n_textures; ++t) {
        const lrt_texture_desc &x = d->textures[t];
        LRT_REQUIRE(x.kind >= 0 && x.kind <= 4, "texture kind must be 0 (bitmap) .. 4 (polarlines)");
        LRT_REQUIRE(x.kind != 0 || (x.width > 0 && x.height > 0 && x.rgb && x.wrap >= 0 && x.wrap <= 2), "bad bitmap texture");
    }
    LRT_REQUIRE(d->n_spheres >= 0 && (d->n_spheres == 0 || d->spheres), "bad sphere table");
    for (int k = 0; k < d->n_spheres; ++k) {
        const float *q = d->spheres + 32 * (size_t) k;
        LRT_REQUIRE((q[6] == 1.f || q[3] > 0.f) && q[4] >= 0.f && q[4] < (float) d->n_shapes && q[4] == (float) (int) q[4], "bad sphere (radius / shape id)");
        LRT_REQUIRE(q[6] == 0.f || (q[6] == 1.f && (q[0] != 0.f || q[1] != 0.f || q[2] != 0.f)), "bad disk (kind / frame normal)");
    }
    int env = -1;
    for (int i = 0; i < d->n_emitters; ++i) {
        const float kind = d->emitters[16 * i + 13];
        LRT_REQUIRE(kind == 0.f || kind == 1.f || kind == 2.f || kind == 3.f, "emitter kind must be 0 (rectangle), 1 (mesh), 2 (constant environment) or 3 (disk)");
        if (kind == 2.f) { LRT_REQUIRE(env < 0, "Only one environment emitter can be specified per scene."); env = i; }
        else LRT_REQUIRE(d->emitters[16 * i + 3] >= 0.f && d->emitters[16 * i + 3] < (float) d->n_shapes, "emitter shape id out of range");
    }
    for (int b = 0; b < d->n_bsdfs; ++b) {
        const float tid = d->bsdfs[16 * b + 15];
        LRT_REQUIRE(tid >= 0.f && tid <= (float) d->n_textures && tid == (float) (int) tid, "bsdf texture id out of range");
    }
    auto t0 = std::chrono::steady_clock::now();
    std::vector<BvhNode> nodes; std::vector<BvhTri> tris; BvhBuildStats st;
    build_bvh8(d->vertices, d->indices, d->tri_shape, d->n_triangles, nodes, tris, &st);
    // per-triangle vertex attributes in BVH order; the triangle record's pad word carries the flags
    std::vector<float> attr;
    if (d->shape_flags) {
        bool any = false;
        // first triangle of every shape (rectangles are the quads (0,1,2),(0,2,3) of their four vertices)
        std::vector<int64_t> first((size_t) d->n_shapes, d->n_triangles);
        for (int64_t t = d->n_triangles - 1; t >= 0; --t) first[d->tri_shape[t]] = t;
        for (size_t i = 0; i < tris.size(); ++i) {
            const int sf = d->shape_flags[tris[i].shape];
            int fl = sf & ((d->vertex_normals ? 1 : 0) | (d->vertex_uvs ? 2 : 0));
            if (sf & 4) {   // rectangle: the frame comes from the triangle's edges; its uvs are only needed by textures
                fl = (tris[i].prim - first[tris[i].shape]) == 0 ? 4 : 8;
                if (d->n_textures > 0 && d->vertex_uvs) fl |= 2;
            }
            memcpy(&tris[i].pad, &fl, 4);
            any = any || (fl & 3) != 0;
        }
        if (any) {
            attr.assign(tris.size() * 16, 0.f);
            for (size_t i = 0; i < tris.size(); ++i) {
                int fl; memcpy(&fl, &tris[i].pad, 4);
                const int32_t *f = d->indices + 3 * (size_t) tris[i].prim; float *o = attr.data() + 16 * i;
                if (fl & 1) for (int k = 0; k < 3; ++k) for (int a = 0; a < 3; ++a) o[3 * k + a] = d->vertex_normals[3 * (size_t) f[k] + a];
                if (fl & 2) for (int k = 0; k < 3; ++k) for (int a = 0; a < 2; ++a) o[9 + 2 * k + a] = d->vertex_uvs[2 * (size_t) f[k] + a];
            }
        }
    }
    std::vector<int32_t> tri_vidx(tris.size() * 3);
    for (size_t i = 0; i < tris.size(); ++i) for (int k = 0; k < 3; ++k) tri_vidx[3 * i + k] = d->indices[3 * (size_t) tris[i].prim + k];
    lrt_scene *s = new lrt_scene();
    s->level_begin = st.level_begin; s->n_vertices = d->n_vertices;
    s->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    s->n_nodes = (int64_t) nodes.size(); s->n_tris = (int64_t) tris.size(); s->bvh_depth = st.max_depth;
    s->n_shapes = d->n_shapes; s->n_bsdfs = d->n_bsdfs; s->n_emitters = d->n_emitters;
    for (int i = 0; i < d->n_emitters; ++i) if (d->emitters[16 * i + 14] == 1.f) s->has_directional = true;
    for (int i = 0; i < d->n_emitters; ++i)
        if (d->emitters[16 * i + 14] == 1.f && d->emitters[16 * i + 13] == 2.f) { set_error("a `directionalarea` emitter needs a shape"); delete s; return LRT_EINVAL; }
    // analytic shapes: a disk's record carries to_world * Normal3f(0, 0, 1); normalised here as disk.cpp:122 does (float32, x * rsqrt)
    std::vector<float> sph(d->spheres, d->spheres + (size_t) d->n_spheres * 32);
    for (int k = 0; k < d->n_spheres; ++k) {
        float *q = sph.data() + 32 * (size_t) k;
        if (q[6] != 1.f) continue;
        const float inv = 1.f / std::sqrt((q[0] * q[0] + q[1] * q[1]) + q[2] * q[2]);
        q[0] *= inv; q[1] *= inv; q[2] *= inv;
    }
    // emitters: 16-float rows -> 5 x float4 with the rectangle's normal and 1/area (rectangle.cpp:115-123)
    std::vector<float> em((size_t) d->n_emitters * 20, 0.f);
    std::vector<float> em_tris;     // 24 floats per triangle of a mesh emitter
    for (int i = 0; i < d->n_emitters; ++i) {
        const float *r = d->emitters + 16 * i; float *o = em.data() + 20 * (size_t) i;
        o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = r[3];
        o[15] = r[14] == 1.f ? 1.f : 0.f;                  // `directionalarea` (directionalarea.cpp): emits along the normal only
        if (r[13] == 2.f) { o[4] = r[4]; o[5] = r[5]; o[6] = r[6]; o[7] = r[7]; o[19] = 2.f; continue; }   // environment: bounding sphere
        if (r[13] == 1.f) {   // area light on a general mesh: Mesh::build_pmf (src/render/mesh.cpp:430-470) + DiscreteDistribution
            const int shape = (int) r[3];
            const bool has_n = d->vertex_normals && d->shape_flags && (d->shape_flags[shape] & 1);
            const int first = (int) (em_tris.size() / 24); int count = 0; double sum = 0.0;
            for (int64_t t = 0; t < d->n_triangles; ++t) {
                if (d->tri_shape[t] != shape) continue;
                const int32_t *f = d->indices + 3 * t;
                const float *p0 = d->vertices + 3 * (size_t) f[0], *p1 = d->vertices + 3 * (size_t) f[1], *p2 = d->vertices + 3 * (size_t) f[2];
                const float a[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] }, b[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
                const float cx = a[1] * b[2] - a[2] * b[1], cy = a[2] * b[0] - a[0] * b[2], cz = a[0] * b[1] - a[1] * b[0];
                const float area = .5f * std::sqrt((cx * cx + cy * cy) + cz * cz);
                sum += (double) area;
                float rec[24] = { p0[0], p0[1], p0[2], (float) sum, a[0], a[1], a[2], area, b[0], b[1], b[2], has_n ? 1.f : 0.f };
                for (int k = 0; k < 3 && has_n; ++k) for (int c = 0; c < 3; ++c) rec[12 + 4 * k + c] = d->vertex_normals[3 * (size_t) f[k] + c];
                em_tris.insert(em_tris.end(), rec, rec + 24); count++;
            }
            LRT_REQUIRE(count > 0 && sum > 0.0, "a mesh emitter needs triangles with a non-zero area");
            const float fsum = em_tris[em_tris.size() - 24 + 3];      // m_sum = the last cdf entry (float)
            memcpy(o + 4, &first, 4); memcpy(o + 5, &count, 4); o[6] = fsum; o[7] = 1.f / fsum; o[19] = 1.f;
            continue;
        }
        o[4] = r[4]; o[5] = r[5]; o[6] = r[6];
        for (int k = 0; k < 3; ++k) { o[8 + k] = r[7 + k]; o[12 + k] = r[10 + k]; }
        if (r[13] == 3.f) {   // on a `disk`: centre, dp_du, dp_dv = translation and the first two columns of to_world; the disk's own frame normal
            const float *q = nullptr;
            for (int k = 0; k < d->n_spheres; ++k) if (sph[32 * (size_t) k + 6] == 1.f && sph[32 * (size_t) k + 4] == r[3]) q = sph.data() + 32 * (size_t) k;
            LRT_REQUIRE(q != nullptr, "emitter kind 3 needs a disk with its shape id");
            o[16] = q[0]; o[17] = q[1]; o[18] = q[2];
            // Disk::surface_area (disk.cpp:163-166): pi |dp_du| h, h = the height of dp_dv over dp_du
            const float du = std::sqrt((r[7] * r[7] + r[8] * r[8]) + r[9] * r[9]), dv = std::sqrt((r[10] * r[10] + r[11] * r[11]) + r[12] * r[12]);
            const float fs[3] = { r[7] * (1.f / du), r[8] * (1.f / du), r[9] * (1.f / du) }, idv = 1.f / dv;
            const float ft[3] = { r[10] * idv * dv, r[11] * idv * dv, r[12] * idv * dv };
            const float dt = (ft[0] * fs[0] + ft[1] * fs[1]) + ft[2] * fs[2];
            const float hh = std::sqrt(dv * dv - dt * dt);
            o[7] = 1.f / (3.14159265358979323846f * du * hh); o[19] = 3.f;
            continue;
        }
        float ex[3] = { r[7], r[8], r[9] }, ey[3] = { r[10], r[11], r[12] };
        float cx = ex[1] * ey[2] - ex[2] * ey[1], cy = ex[2] * ey[0] - ex[0] * ey[2], cz = ex[0] * ey[1] - ex[1] * ey[0];
        float len = std::sqrt(cx * cx + cy * cy + cz * cz);
        o[16] = cx / len; o[17] = cy / len; o[18] = cz / len;
        o[7] = 1.f / (4.f * len);
    }
    int rc = LRT_OK;
    if ((rc = upload(&s->d_nodes, nodes.data(), nodes.size())) || (rc = upload(&s->d_tris, tris.data(), tris.size())) ||
        (rc = upload(&s->d_shape_bsdf, d->shape_bsdf, (size_t) d->n_shapes)) ||
        (rc = upload(&s->d_shape_emitter, d->shape_emitter, (size_t) d->n_shapes)) ||
        (rc = upload(&s->d_bsdfs, d->bsdfs, (size_t) d->n_bsdfs * 16)) || (rc = upload(&s->d_emitters, em.data(), em.size())) ||
        (!em_tris.empty() && (rc = upload(&s->d_em_tris, em_tris.data(), em_tris.size())))) {
        lrt_scene_destroy(s); return rc;
    }
    if ((rc = upload(&s->d_tri_vidx, tri_vidx.data(), tri_vidx.size()))) { lrt_scene_destroy(s); return rc; }
    if (cudaMalloc(&s->d_node_box, std::max<size_t>(nodes.size(), 1) * 6 * sizeof(float)) != cudaSuccess) { lrt_scene_destroy(s); return cuda_fail(cudaGetLastError(), "cudaMalloc"); }
    if (!attr.empty() && (rc = upload(&s->d_tri_attr, attr.data(), attr.size()))) { lrt_scene_destroy(s); return rc; }
    if (d->n_textures) {
        auto texels_of = [&](int t) { return d->textures[t].kind == 0 ? (size_t) d->textures[t].width * d->textures[t].height * 3 : (size_t) 0; };
        size_t total = 0;
        for (int t = 0; t < d->n_textures; ++t) total += texels_of(t);
        std::vector<float> texels; texels.reserve(total);
        for (int t = 0; t < d->n_textures; ++t) if (texels_of(t)) texels.insert(texels.end(), d->textures[t].rgb, d->textures[t].rgb + texels_of(t));
        if ((rc = upload(&s->d_tex_data, texels.data(), texels.size()))) { lrt_scene_destroy(s); return rc; }
        std::vector<TexView> tv(d->n_textures); size_t off = 0;
        for (int t = 0; t < d->n_textures; ++t) {
            const lrt_texture_desc &x = d->textures[t];
            tv[t].data = (const float *) s->d_tex_data + off; tv[t].w = x.width; tv[t].h = x.height;
            tv[t].wrap = x.wrap; tv[t].kind = x.kind;
            for (int k = 0; k < 6; ++k) tv[t].to_uv[k] = x.to_uv[k];
            for (int k = 0; k < 3; ++k) { tv[t].color0[k] = x.color0[k]; tv[t].color1[k] = x.color1[k]; }
            tv[t].linewidth = x.linewidth; tv[t].spokes = x.spokes;
            off += texels_of(t);
        }
        if ((rc = upload(&s->d_textures, tv.data(), tv.size()))) { lrt_scene_destroy(s); return rc; }
        s->n_textures = d->n_textures;
    }
    if (d->n_spheres && (rc = upload(&s->d_spheres, sph.data(), sph.size()))) { lrt_scene_destroy(s); return rc; }
    if (cudaMalloc(&s->d_counter, (kCounterRing + 8) * sizeof(unsigned long long)) != cudaSuccess) { lrt_scene_destroy(s); return cuda_fail(cudaGetLastError(), "cudaMalloc"); }
    s->view.bvh.nodes = (const uint4 *) s->d_nodes; s->view.bvh.tris = (const float4 *) s->d_tris;
    s->view.bvh.kb = 0x4B000000u; s->view.bvh.km = 0xfffffff8u;
    s->view.shape_bsdf = (const int *) s->d_shape_bsdf; s->view.shape_emitter = (const int *) s->d_shape_emitter;
    s->view.bsdfs = (const float4 *) s->d_bsdfs; s->view.emitters = (const float4 *) s->d_emitters;
    s->view.n_emitters = d->n_emitters;
    s->view.tri_attr = (const float4 *) s->d_tri_attr; s->view.textures = (const TexView *) s->d_textures;
    s->view.em_tris = (const float4 *) s->d_em_tris;
    s->view.spheres = (const float4 *) s->d_spheres; s->view.n_spheres = d->n_spheres; s->view.n_tris = (int) s->n_tris;
    s->view.env = env;
    *out = s;
    return LRT_OK;
}

// parameters_changed() for the differentiable scene parameters: new BSDF rows (same count) and / or emitter radiances
int lrt_scene_set_params(lrt_scene *s, const float *bsdfs, const float *emitter_radiance3) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s, "null scene");
    if (bsdfs && s->n_bsdfs) LRT_CUDA(cudaMemcpy(s->d_bsdfs, bsdfs, (size_t) s->n_bsdfs * 64, cudaMemcpyHostToDevice));
    if (emitter_radiance3)
        for (int i = 0; i < s->n_emitters; ++i)   // radiance is the first 3 floats of the 20-float device record
            LRT_CUDA(cudaMemcpy((float *) s->d_emitters + 20 * (size_t) i, emitter_radiance3 + 3 * i, 12, cudaMemcpyHostToDevice));
    return LRT_OK;
}

void lrt_scene_destroy(lrt_scene *s) {
    if (!s) return;
    void *p[] = { s->d_nodes, s->d_tris, s->d_shape_bsdf, s->d_shape_emitter, s->d_bsdfs, s->d_emitters, s->d_counter,
                  s->d_tri_attr, s->d_tex_data, s->d_textures, s->d_tri_vidx, s->d_node_box, s->d_em_tris, s->d_spheres };
    for (void *q : p) if (q) cudaFree(q);
    delete s;
}

int lrt_scene_info(const lrt_scene *s, int64_t *n_nodes, int64_t *n_tris, int32_t *bvh_depth, double *build_seconds) {
    LRT_REQUIRE(s, "null scene");
    if (n_nodes) *n_nodes = s->n_nodes;
    if (n_tris) *n_tris = s->n_tris;
    if (bvh_depth) *bvh_depth = s->bvh_depth;
    if (build_seconds) *build_seconds = s->build_seconds;
    return LRT_OK;
}

int lrt_render(const lrt_scene *sc, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
               float *accum_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && sensor && film && rp && accum_dev, "null argument");
    LRT_REQUIRE(rp->rng_mode == 0 || rp->rng_mode == 1, "rng_mode must be 0 (per-sample streams) or 1 (scalar convention)");
    LRT_REQUIRE(!sc->has_directional, "`directionalarea` emitters are built for the particle tracer only (lrt_ptrace)");
    LRT_REQUIRE((reinterpret_cast<uintptr_t>(accum_dev) & 15) == 0, "accum_dev must be 16-byte aligned (the tiles are flushed with 128-bit reductions)");
    if (rp->rng_mode == 1) return launch_render_scalar(sc, const_cast<lrt_sensor *>(sensor), film, rp, accum_dev, (cudaStream_t) stream);
    return launch_render(sc, const_cast<lrt_sensor *>(sensor), film, rp, accum_dev, nullptr, nullptr, nullptr, (cudaStream_t) stream);
}

int lrt_render_backward(const lrt_scene *sc, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                        const float *grad_storage_dev, float *grad_bsdfs_dev, float *grad_emitters_dev, void *stream) {
    LRT_REQUIRE(!sc || !sc->has_directional, "`directionalarea` emitters are built for the particle tracer only (lrt_ptrace)");
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && sensor && film && rp && grad_storage_dev && grad_bsdfs_dev && grad_emitters_dev, "null argument");
    LRT_REQUIRE(rp->rng_mode == 0, "the adjoint follows the JIT sampler convention (rng_mode 0)");
    return launch_render(sc, const_cast<lrt_sensor *>(sensor), film, rp, nullptr, grad_storage_dev, grad_bsdfs_dev, grad_emitters_dev,
                         (cudaStream_t) stream);
}

int lrt_render_stats(const lrt_scene *sc, uint64_t *out7) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && out7, "null argument");
    LRT_CUDA(cudaMemcpy(out7, sc->d_counter + kCounterRing, 7 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return LRT_OK;
}

// mi.render() for host callers: accumulate + develop; rgb_out [size_y, size_x, 3] on the host
int lrt_render_host(const lrt_scene *sc, const lrt_sensor *sensor, const lrt_film_desc *film, const lrt_render_params *rp,
                    float *rgb_out) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(sc && sensor && film && rp && rgb_out, "null argument");
    LRT_REQUIRE(film->border == 0 && film->channels == 4, "render_host needs a borderless R,G,B,W film");
    size_t npix = (size_t) film->size_x * film->size_y;
    float *acc = nullptr, *rgb = nullptr;
    LRT_CUDA(cudaMalloc(&acc, npix * 16)); 
    if (cudaMalloc(&rgb, npix * 12) != cudaSuccess) { cudaFree(acc); return cuda_fail(cudaGetLastError(), "cudaMalloc"); }
    cudaMemsetAsync(acc, 0, npix * 16, nullptr);
    int rc = lrt_render(sc, sensor, film, rp, acc, nullptr);
    if (rc == LRT_OK) rc = lrt_film_develop_rgb((int64_t) npix, acc, rgb, nullptr);
    if (rc == LRT_OK) { cudaError_t e = cudaMemcpy(rgb_out, rgb, npix * 12, cudaMemcpyDeviceToHost); if (e != cudaSuccess) rc = cuda_fail(e, "copy back"); }
    cudaFree(acc); cudaFree(rgb);
    return rc;
}

}  // extern "C"
