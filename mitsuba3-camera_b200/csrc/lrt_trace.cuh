// lensrt — device traversal of the 8-wide quantised BVH (lrt_bvh.h) + Moeller-Trumbore triangle test.
// Replaces Scene::ray_intersect / ray_test (src/render/scene.cpp:181-210) and
// Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:432-452; same formulation, same acceptance tests).
// One thread = one ray; nodes and triangles are fetched as 128-bit read-only loads; the traversal stack
// lives in local memory (L1-resident) and stores (t_near, reference) so popped entries can be culled.
#pragma once
#include "lrt_math.cuh"
#include "lrt_bvh.h"

namespace lrt {

struct RayHit { float t, u, v; int tri; };   // tri = index into the reordered triangle array, -1 = miss
// `work` accumulates the algorithmic traffic of a traversal: nodes fetched (low 16 bits) and triangles tested
// (high 16 bits); the render kernel sums it into its statistics (bench.py's roofline `achieved`).

LRT_D float q2f(uint32_t w, int byte) { return (float) ((w >> (8 * byte)) & 0xffu); }

template <bool ANY_HIT>
LRT_D RayHit bvh_intersect(const BvhView &bvh, float3 o, float3 d, float tmax, unsigned &work) {
    RayHit hit; hit.t = inf_f(); hit.u = hit.v = 0.f; hit.tri = -1;
    // guard against 0 * inf in the slab test
    const float tiny = 1e-30f;
    float3 dd = make_float3(fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y,
                            fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
    float3 idir = make_float3(1.f / dd.x, 1.f / dd.y, 1.f / dd.z);
    const bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;

    constexpr int kStack = 64;
    uint32_t st_ref[kStack]; float st_t[kStack];
    int sp = 0;
    uint32_t cur = 0;            // bit31 = leaf; leaf: bits 29..28 count, 27..0 first triangle
    float cur_t = 0.f;
    while (true) {
        if (cur_t <= tmax) {
            if (!(cur & 0x80000000u)) {
                work += 1u;
                const uint4 *np = bvh.nodes + 5 * (size_t) cur;
                uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
                float sx = __uint_as_float((n0.w & 0xffu) << 23) * idir.x, sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * idir.y,
                      sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * idir.z;
                float bx = (__uint_as_float(n0.x) - o.x) * idir.x, by = (__uint_as_float(n0.y) - o.y) * idir.y,
                      bz = (__uint_as_float(n0.z) - o.z) * idir.z;
                uint32_t best = 0xffffffffu; float best_t = inf_f();
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    uint32_t meta = ((c < 4 ? n1.z : n1.w) >> (8 * (c & 3))) & 0xffu;
                    if (meta == 0) continue;
                    const int b = c & 3;
                    float lox = q2f(c < 4 ? n2.x : n2.y, b), loy = q2f(c < 4 ? n2.z : n2.w, b), loz = q2f(c < 4 ? n3.x : n3.y, b);
                    float hix = q2f(c < 4 ? n3.z : n3.w, b), hiy = q2f(c < 4 ? n4.x : n4.y, b), hiz = q2f(c < 4 ? n4.z : n4.w, b);
                    float t0x = fmaf(nx ? hix : lox, sx, bx), t1x = fmaf(nx ? lox : hix, sx, bx);
                    float t0y = fmaf(ny ? hiy : loy, sy, by), t1y = fmaf(ny ? loy : hiy, sy, by);
                    float t0z = fmaf(nz ? hiz : loz, sz, bz), t1z = fmaf(nz ? loz : hiz, sz, bz);
                    float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
                    float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax)) * 1.0000005f;
                    if (tn <= tf) {
                        uint32_t ref = (meta & 0x80u) ? (n1.x + (meta & 0x7fu))
                                                      : (0x80000000u | ((meta >> 5) << 28) | (n1.y + (meta & 0x1fu)));
                        if (tn < best_t) {
                            if (best != 0xffffffffu && sp < kStack) { st_ref[sp] = best; st_t[sp] = best_t; sp++; }
                            best = ref; best_t = tn;
                        } else if (sp < kStack) { st_ref[sp] = ref; st_t[sp] = tn; sp++; }
                    }
                }
                if (best != 0xffffffffu) { cur = best; cur_t = best_t; continue; }
            } else {
                uint32_t first = cur & 0x0fffffffu, count = (cur >> 28) & 3u;
                work += count << 16;
                for (uint32_t k = 0; k < count; ++k) {
                    const float4 *tp = bvh.tris + 3 * (size_t) (first + k);
                    float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                    float3 p0 = make_float3(a.x, a.y, a.z), p1 = make_float3(a.w, b.x, b.y), p2 = make_float3(b.z, b.w, c.x);
                    // mesh.h:432-452
                    float3 e1 = p1 - p0, e2 = p2 - p0;
                    float3 pvec = cross(d, e2);
                    float inv_det = 1.f / dot(e1, pvec);
                    float3 tvec = o - p0;
                    float u = dot(tvec, pvec) * inv_det;
                    float3 qvec = cross(tvec, e1);
                    float v = dot(d, qvec) * inv_det;
                    float t = dot(e2, qvec) * inv_det;
                    bool ok = u >= 0.f && u <= 1.f && v >= 0.f && u + v <= 1.f && t >= 0.f && t <= tmax;
                    if (ok && (t < hit.t || (t == hit.t && (int) (first + k) < hit.tri))) {
                        hit.t = t; hit.u = u; hit.v = v; hit.tri = (int) (first + k); tmax = t;
                        if (ANY_HIT) return hit;
                    }
                }
            }
        }
        if (sp == 0) break;
        --sp; cur = st_ref[sp]; cur_t = st_t[sp];
    }
    return hit;
}

}  // namespace lrt
