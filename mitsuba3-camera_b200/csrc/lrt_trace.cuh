// lensrt — device traversal of the 8-wide quantised BVH (lrt_bvh.h) + Moeller-Trumbore triangle test.
// Replaces Scene::ray_intersect / ray_test (src/render/scene.cpp:181-210) and
// Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:432-452; same formulation, same acceptance tests).
//
// One thread = one ray. Closest-hit and any-hit (shadow) rays share ONE code path (`any_hit` is a run-time
// flag) so that the render kernel has a single traversal site: lanes tracing shadow rays and lanes tracing
// path segments run the same instructions, and the hot loop fits the instruction cache.
// Nodes and triangles are fetched as 128-bit read-only loads. Child bounds are decoded without I2F: a PRMT
// splices the byte into the mantissa of 2^23 and one FFMA per bound maps it to ray space (the -2^23 bias is
// folded into the per-node offset; the builder keeps one full quantisation step of slack on every side, more
// than the <= 1/2 step this costs in precision). The traversal stack (t_near, reference) lives in local memory.
#pragma once
#include "lrt_math.cuh"
#include "lrt_bvh.h"

namespace lrt {

struct RayHit { float t, u, v; int tri; };   // tri = index into the reordered triangle array, -1 = miss
// `work` accumulates the algorithmic traffic of a traversal: nodes fetched (low 16 bits) and triangles tested
// (high 16 bits); the render kernel sums it into its statistics (bench.py's roofline `achieved`).

// 0x4B0000qq as float = 8388608 + q
LRT_D float qbias(uint32_t w, uint32_t sel) { return __uint_as_float(__byte_perm(w, 0x4B000000u, sel)); }

LRT_D RayHit bvh_intersect(const BvhView &bvh, float3 o, float3 d, float tmax, bool any_hit, unsigned &work) {
    RayHit hit; hit.t = inf_f(); hit.u = hit.v = 0.f; hit.tri = -1;
    const float tiny = 1e-30f;   // guard against 0 * inf in the slab test
    float3 idir = make_float3(1.f / (fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x), 1.f / (fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y),
                              1.f / (fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z));
    const bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;

    constexpr int kStack = 64;
    constexpr uint32_t kDone = 0x7fffffffu;
    uint32_t st_ref[kStack]; float st_t[kStack];
    int sp = 0;
    uint32_t cur = 0;            // bit31 = leaf; leaf: bits 29..28 count, 27..0 first triangle
    while (cur != kDone) {
        // ---- inner nodes: descend to the nearest child until a leaf (or nothing) is reached
        while (!(cur & 0x80000000u)) {
            work += 1u;
            const uint4 *np = bvh.nodes + 5 * (size_t) cur;
            const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            const float sx = __uint_as_float((n0.w & 0xffu) << 23) * idir.x, sy = __uint_as_float(((n0.w >> 8) & 0xffu) << 23) * idir.y,
                        sz = __uint_as_float(((n0.w >> 16) & 0xffu) << 23) * idir.z;
            const float bx = fmaf(-8388608.f, sx, (__uint_as_float(n0.x) - o.x) * idir.x),
                        by = fmaf(-8388608.f, sy, (__uint_as_float(n0.y) - o.y) * idir.y),
                        bz = fmaf(-8388608.f, sz, (__uint_as_float(n0.z) - o.z) * idir.z);
            // near / far bound words per axis, children 0-3 (a) and 4-7 (b)
            const uint32_t nxa = nx ? n3.z : n2.x, nxb = nx ? n3.w : n2.y, fxa = nx ? n2.x : n3.z, fxb = nx ? n2.y : n3.w;
            const uint32_t nya = ny ? n4.x : n2.z, nyb = ny ? n4.y : n2.w, fya = ny ? n2.z : n4.x, fyb = ny ? n2.w : n4.y;
            const uint32_t nza = nz ? n4.z : n3.x, nzb = nz ? n4.w : n3.y, fza = nz ? n3.x : n4.z, fzb = nz ? n3.y : n4.w;
            uint32_t best = kDone; float best_t = inf_f();
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint32_t meta = ((c < 4 ? n1.z : n1.w) >> (8 * (c & 3))) & 0xffu;
                const uint32_t sel = 0x7440u | (uint32_t) (c & 3);
                const float t0x = fmaf(qbias(c < 4 ? nxa : nxb, sel), sx, bx), t1x = fmaf(qbias(c < 4 ? fxa : fxb, sel), sx, bx);
                const float t0y = fmaf(qbias(c < 4 ? nya : nyb, sel), sy, by), t1y = fmaf(qbias(c < 4 ? fya : fyb, sel), sy, by);
                const float t0z = fmaf(qbias(c < 4 ? nza : nzb, sel), sz, bz), t1z = fmaf(qbias(c < 4 ? fza : fzb, sel), sz, bz);
                const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
                const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
                if (meta != 0 && tn <= tf) {
                    const uint32_t ref = (meta & 0x80u) ? (n1.x + (meta & 0x7fu))
                                                        : (0x80000000u | ((meta >> 5) << 28) | (n1.y + (meta & 0x1fu)));
                    const bool closer = tn < best_t;
                    const uint32_t push_ref = closer ? best : ref; const float push_t = closer ? best_t : tn;
                    if (closer) { best = ref; best_t = tn; }
                    if (push_ref != kDone && sp < kStack) { st_ref[sp] = push_ref; st_t[sp] = push_t; sp++; }
                }
            }
            cur = best;
            if (cur == kDone) {   // nothing hit: pop the next entry that can still matter
                while (sp > 0) { --sp; if (st_t[sp] <= tmax) { cur = st_ref[sp]; break; } }
                if (cur == kDone) break;
            }
        }
        if (cur == kDone) break;
        // ---- leaf: up to 3 triangles
        {
            const uint32_t first = cur & 0x0fffffffu, count = (cur >> 28) & 3u;
            work += count << 16;
            for (uint32_t k = 0; k < count; ++k) {
                const float4 *tp = bvh.tris + 3 * (size_t) (first + k);
                const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                const float3 p0 = make_float3(a.x, a.y, a.z), p1 = make_float3(a.w, b.x, b.y), p2 = make_float3(b.z, b.w, c.x);
                // mesh.h:432-452
                const float3 e1 = p1 - p0, e2 = p2 - p0;
                const float3 pvec = cross(d, e2);
                const float inv_det = 1.f / dot(e1, pvec);
                const float3 tvec = o - p0;
                const float u = dot(tvec, pvec) * inv_det;
                const float3 qvec = cross(tvec, e1);
                const float v = dot(d, qvec) * inv_det;
                const float t = dot(e2, qvec) * inv_det;
                const bool ok = u >= 0.f && u <= 1.f && v >= 0.f && u + v <= 1.f && t >= 0.f && t <= tmax;
                if (ok && (t < hit.t || (t == hit.t && (int) (first + k) < hit.tri))) {
                    hit.t = t; hit.u = u; hit.v = v; hit.tri = (int) (first + k); tmax = t;
                }
            }
            if (any_hit && hit.tri >= 0) return hit;
            cur = kDone;
            while (sp > 0) { --sp; if (st_t[sp] <= tmax) { cur = st_ref[sp]; break; } }
        }
    }
    return hit;
}

}  // namespace lrt
