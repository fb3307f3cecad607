// lensrt — mesh-lens parity mode, device side: BVH REFIT after the vertices moved, and a ray trace
// through refractive meshes.
//
// The notebooks (notebooks/scripts/lens_design/lens.py:343-565, geometry.py:46-300) model every lens surface as a triangle
// mesh whose vertex z follows the sag function; each optimisation step moves the vertices (params['lensN.vertex_positions'],
// lens.py:67-85,210-237) and the renderer rebuilds its acceleration structure (Mesh::parameters_changed,
// src/render/mesh.cpp:85-162 -> Scene::parameters_changed -> OptiX / Embree / kd-tree rebuild). Here the topology of the
// 8-wide BVH is kept and only its boxes are refitted on the device: one kernel rewrites the triangle records (and the vertex
// normals), one small kernel per BFS level — bottom-up — recomputes every node's grid (origin, per-axis exponent) and its
// children's quantised bounds exactly as the host builder does (lrt_bvh_build.cu).
//
// lrt_scene_trace_refractive follows rays through such meshes: closest hit -> refraction about the interpolated vertex
// normal with the hit shape's relative index (fresnel / refract of include/mitsuba/render/fresnel.h:35-84,311-315 as the
// `dispersive` / `dielectric` BSDF's transmission lobe applies them, src/bsdfs/dispersive_dielectric.cpp:250-431) -> next
// surface; shapes with eta = 0 absorb (the notebooks' black baffles). What is left is the ray that leaves the lens.
#include "lrt_scene.h"
#include <cmath>

namespace lrt {

__global__ void __launch_bounds__(256) k_refit_tris(int64_t nt, const int3 *__restrict__ vidx, const float *__restrict__ verts,
                                                     const float *__restrict__ normals, float4 *__restrict__ tris, float4 *__restrict__ attr) {
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < nt; i += (int64_t) gridDim.x * blockDim.x) {
        const int3 f = vidx[i];
        const float *a = verts + 3 * (size_t) f.x, *b = verts + 3 * (size_t) f.y, *c = verts + 3 * (size_t) f.z;
        float4 *t = tris + 3 * i;
        const float4 t2 = t[2];
        t[0] = make_float4(a[0], a[1], a[2], b[0]);
        t[1] = make_float4(b[1], b[2], c[0], c[1]);
        t[2] = make_float4(c[2], t2.y, t2.z, t2.w);
        if (normals && attr && (__float_as_int(t2.w) & 1)) {
            const float *n0 = normals + 3 * (size_t) f.x, *n1 = normals + 3 * (size_t) f.y, *n2 = normals + 3 * (size_t) f.z;
            float4 *q = attr + 4 * i;
            const float4 q2 = q[2];
            q[0] = make_float4(n0[0], n0[1], n0[2], n1[0]);
            q[1] = make_float4(n1[1], n1[2], n2[0], n2[1]);
            q[2] = make_float4(n2[2], q2.y, q2.z, q2.w);
        }
    }
}

// smallest biased exponent e with 253 * 2^(e-127) >= extent (lrt_bvh_build.cu exponent_for)
__device__ inline int refit_exponent(float extent) {
    if (!(extent > 0.f)) return 1;
    int e; frexpf(extent / 253.f, &e);
    return min(254, max(1, e + 127));
}

__global__ void __launch_bounds__(128) k_refit_level(int first, int last, uint4 *__restrict__ nodes, const float4 *__restrict__ tris,
                                                      float *__restrict__ node_box) {
    const int n = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= last) return;
    uint4 *np = nodes + 5 * (size_t) n;
    const uint4 n1 = np[1];
    float lo[8][3], hi[8][3];
    float blo[3] = { INFINITY, INFINITY, INFINITY }, bhi[3] = { -INFINITY, -INFINITY, -INFINITY };
    unsigned used = 0;
    for (int c = 0; c < 8; ++c) {
        const unsigned meta = ((c < 4 ? n1.z : n1.w) >> (8 * (c & 3))) & 0xffu;
        if (meta == 0) continue;
        used |= 1u << c;
        for (int a = 0; a < 3; ++a) { lo[c][a] = INFINITY; hi[c][a] = -INFINITY; }
        if (meta & 0x80u) {
            const float *cb = node_box + 6 * (size_t) (n1.x + (meta & 0x7fu));
            for (int a = 0; a < 3; ++a) { lo[c][a] = cb[a]; hi[c][a] = cb[3 + a]; }
        } else {
            const unsigned cnt = meta >> 5, off = meta & 0x1fu;
            for (unsigned k = 0; k < cnt; ++k) {
                const float4 *t = tris + 3 * (size_t) (n1.y + off + k);
                const float4 a = t[0], b = t[1], cc = t[2];
                const float p[3][3] = { { a.x, a.y, a.z }, { a.w, b.x, b.y }, { b.z, b.w, cc.x } };
                for (int v = 0; v < 3; ++v) for (int ax = 0; ax < 3; ++ax) { lo[c][ax] = fminf(lo[c][ax], p[v][ax]); hi[c][ax] = fmaxf(hi[c][ax], p[v][ax]); }
            }
        }
        for (int a = 0; a < 3; ++a) { blo[a] = fminf(blo[a], lo[c][a]); bhi[a] = fmaxf(bhi[a], hi[c][a]); }
    }
    float *nb = node_box + 6 * (size_t) n;
    for (int a = 0; a < 3; ++a) { nb[a] = blo[a]; nb[3 + a] = bhi[a]; }
    if (!used) return;
    int ex[3]; float sc[3], org[3];
    for (int a = 0; a < 3; ++a) { ex[a] = refit_exponent(bhi[a] - blo[a]); sc[a] = ldexpf(1.f, ex[a] - 127); org[a] = blo[a] - sc[a]; }
    unsigned char q[6][8];
    for (int c = 0; c < 8; ++c) {
        for (int a = 0; a < 3; ++a) {
            if (!(used & (1u << c))) { q[a][c] = 255; q[3 + a][c] = 0; continue; }       // inverted box: never hit
            const float l = floorf((lo[c][a] - org[a]) / sc[a]), h = ceilf((hi[c][a] - org[a]) / sc[a]);
            q[a][c] = (unsigned char) fminf(255.f, fmaxf(0.f, l - 1.f));
            q[3 + a][c] = (unsigned char) fminf(255.f, fmaxf(0.f, h + 1.f));
        }
    }
    auto pack = [&](const unsigned char *b) { return (unsigned) b[0] | ((unsigned) b[1] << 8) | ((unsigned) b[2] << 16) | ((unsigned) b[3] << 24); };
    np[0] = make_uint4(__float_as_uint(org[0]), __float_as_uint(org[1]), __float_as_uint(org[2]),
                       (unsigned) ex[0] | ((unsigned) ex[1] << 8) | ((unsigned) ex[2] << 16));
    // words 2-4: qlo_x[8] qlo_y[8] | qlo_z[8] qhi_x[8] | qhi_y[8] qhi_z[8]  (lrt_bvh.h)
    np[2] = make_uint4(pack(q[0]), pack(q[0] + 4), pack(q[1]), pack(q[1] + 4));
    np[3] = make_uint4(pack(q[2]), pack(q[2] + 4), pack(q[3]), pack(q[3] + 4));
    np[4] = make_uint4(pack(q[4]), pack(q[4] + 4), pack(q[5]), pack(q[5] + 4));
}

// rays through refractive meshes: od6 -> out8 = o3, d3, alive (1/0), weight (product of the Fresnel transmittances)
__global__ void __launch_bounds__(128) k_trace_refractive(SceneView sc, int64_t n, const float *__restrict__ od6, const float *__restrict__ shape_eta,
                                                           int max_interactions, float *__restrict__ out8) {
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float *r = od6 + 6 * i;
        float3 o = make_float3(r[0], r[1], r[2]), d = normalize(make_float3(r[3], r[4], r[5]));
        float weight = 1.f; bool alive = true;
        for (int k = 0; k < max_interactions; ++k) {
            unsigned work = 0;
            const RayHit h = bvh_intersect(sc.bvh, o, d, kLargest, false, work);
            if (h.tri < 0) break;
            const Surface si = make_surface(sc, h, o, d);
            const float eta = __ldg(shape_eta + si.shape);
            if (!(eta > 0.f)) { alive = false; break; }                    // absorber (baffle)
            float r_i, ct, eit, eti;
            fresnel_dielectric(si.wi.z, eta, r_i, ct, eit, eti);
            if (r_i >= 1.f) { alive = false; break; }                      // total internal reflection
            weight *= 1.f - r_i;
            const float3 wo = make_float3(-eti * si.wi.x, -eti * si.wi.y, ct);   // refract(wi, cos_theta_t, eta_ti)
            d = normalize(si.to_world(wo));
            // the next origin moves ALONG the refracted ray (its line is unchanged). The reference's spawn_ray steps
            // RayEpsilon * (1 + |p|) along the NORMAL (interaction.h:156-162), which displaces a refracted ray sideways by up to
            // ~1e-3 mm in the notebooks' millimetre-scale scenes — the order of the meshing error they report; not reproduced.
            o = si.p + d * (kRayEpsilon * 0.01f * (1.f + fmaxf(fabsf(si.p.x), fmaxf(fabsf(si.p.y), fabsf(si.p.z)))));
        }
        float *q = out8 + 8 * i;
        q[0] = o.x; q[1] = o.y; q[2] = o.z; q[3] = d.x; q[4] = d.y; q[5] = d.z; q[6] = alive ? 1.f : 0.f; q[7] = alive ? weight : 0.f;
    }
}

}  // namespace lrt

using namespace lrt;
extern "C" {

int lrt_scene_refit(lrt_scene *s, const float *vertices_dev, const float *normals_dev, void *stream_) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && vertices_dev, "null argument");
    cudaStream_t stream = (cudaStream_t) stream_;
    if (s->n_tris == 0) return LRT_OK;
    LRT_REQUIRE(!normals_dev || s->d_tri_attr, "the scene was created without vertex normals");
    int64_t g = (s->n_tris + 255) / 256; if (g > 148 * 16) g = 148 * 16;
    k_refit_tris<<<(int) g, 256, 0, stream>>>(s->n_tris, (const int3 *) s->d_tri_vidx, vertices_dev, normals_dev, (float4 *) s->d_tris,
                                             (float4 *) s->d_tri_attr);
    LRT_LAUNCHED();
    for (int l = (int) s->level_begin.size() - 2; l >= 0; --l) {           // deepest level first: children before parents
        const int first = s->level_begin[l], last = s->level_begin[l + 1];
        if (last <= first) continue;
        k_refit_level<<<(last - first + 127) / 128, 128, 0, stream>>>(first, last, (uint4 *) s->d_nodes, (const float4 *) s->d_tris, (float *) s->d_node_box);
        LRT_LAUNCHED();
    }
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

int lrt_scene_trace_refractive(const lrt_scene *s, int64_t n, const float *od6_dev, const float *shape_eta_dev, int32_t max_interactions,
                               float *out8_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && n >= 0 && max_interactions >= 0 && (n == 0 || (od6_dev && shape_eta_dev && out8_dev)), "bad argument");
    if (n == 0) return LRT_OK;
    int64_t g = (n + 127) / 128; if (g > 148 * 32) g = 148 * 32;
    k_trace_refractive<<<(int) g, 128, 0, (cudaStream_t) stream>>>(s->view, n, od6_dev, shape_eta_dev, max_interactions, out8_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

}  // extern "C"
