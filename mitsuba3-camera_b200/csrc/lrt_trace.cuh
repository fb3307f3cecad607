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

// 0x4B0000qq as float = 8388608 + q. `kb` = 0x4B000000 comes from BvhView (data, not a literal), so that the byte selector
// and not the constant becomes PRMT's immediate operand: no selector MOVs in the unrolled node test.
LRT_D float qbias(uint32_t w, uint32_t kb, uint32_t sel) { return __uint_as_float(__byte_perm(w, kb, sel)); }

constexpr int kTravStack = 64;
constexpr uint32_t kTravDone = 0x7fffffffu;

// Resumable traversal state of one ray: `cur` = node / leaf to visit next (bit31 = leaf: bits 29..28 count, 27..0 first
// triangle; kTravDone = finished), the stack (t_near, reference) lives in the caller's local memory.
struct Trav { uint32_t cur; int sp; RayHit hit; float tmax; };

LRT_D void trav_init(Trav &T, float tmax) {
    T.cur = 0; T.sp = 0; T.tmax = tmax;
    T.hit.t = inf_f(); T.hit.u = T.hit.v = 0.f; T.hit.tri = -1;
}

LRT_D float3 trav_idir(float3 d) {
    const float tiny = 1e-30f;   // guard against 0 * inf in the slab test
    return make_float3(1.f / (fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x), 1.f / (fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y),
                       1.f / (fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z));
}

// child slot c of a node (n1 = its second 128-bit word) -> traversal reference (inner node index | leaf encoding)
LRT_D uint32_t trav_child_ref(const uint4 &n1, uint32_t c) {
    const uint32_t meta = ((c < 4u ? n1.z : n1.w) >> (8u * (c & 3u))) & 0xffu;
    return (meta & 0x80u) ? (n1.x + (meta & 0x7fu)) : (0x80000000u | ((meta >> 5) << 28) | (n1.y + (meta & 0x1fu)));
}

// Stack entries are (node index, key) with key = (bits of t_near & ~7) | child slot, one 64-bit word each: children are
// pushed without decoding them; the reference is rebuilt from the node's second word only when an entry is popped.
LRT_D uint32_t trav_pop(const BvhView &bvh, int &sp, const uint2 *st, float tmax) {
    while (sp > 0) {
        --sp;
        const uint2 e = st[sp];
        if (__uint_as_float(e.y & ~7u) <= tmax)
            return trav_child_ref(__ldg(bvh.nodes + 5 * (size_t) (e.x & 0x7fffffffu) + 1), (e.y & 3u) | ((e.x >> 31) << 2));
    }
    return kTravDone;
}

// One inner node: test the 8 quantised child boxes, continue with the nearest hit child, push the others; if nothing is
// hit pop the next candidate. Requires T.cur to be an inner node. Empty child slots carry an inverted box (lo 255,
// hi 0: the builder writes it), so they fail the slab test by themselves and `meta` is not looked at here.
// The children are tested in two rounds of four by a loop that is NOT unrolled: the traversal loop then fits the
// instruction cache next to the leaf test. Keys carry the slot within the round (2 bits); the round is bit 31 of the
// stack entry's node word.
LRT_D void trav_node_step(const BvhView &bvh, float3 o, float3 idir, Trav &T, uint2 *st, unsigned &work) {
    const uint32_t kb = bvh.kb, km = bvh.km;   // 0x4B000000 and ~7 as register operands, see qbias() and the key below
    const bool nx = idir.x < 0.f, ny = idir.y < 0.f, nz = idir.z < 0.f;
    const float tmax = T.tmax;
    work += 1u;
    const uint4 *np = bvh.nodes + 5 * (size_t) T.cur;
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
    // a full stack cannot take this node's (up to 8) pushes: the newest candidates are overwritten and the event is flagged in
    // bit 31 of `work` (the render kernel counts it into its statistics: tests and bench assert the count is zero; a BVH8 over
    // a million triangles keeps < 30 entries)
    if (T.sp > kTravStack - 9) work |= 0x80000000u;
    int sp = min(T.sp, kTravStack - 9);
    uint32_t best = 0xffffffffu, best0 = 0xffffffffu, xh = T.cur;
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
        best0 = best; best = 0xffffffffu;   // after the loop: best0 = nearest of children 0-3, best = nearest of 4-7
        const uint32_t wnx = h ? nxb : nxa, wfx = h ? fxb : fxa, wny = h ? nyb : nya, wfy = h ? fyb : fya, wnz = h ? nzb : nza, wfz = h ? fzb : fza;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t sel = 0x7440u | (uint32_t) c;
            const float t0x = fmaf(qbias(wnx, kb, sel), sx, bx), t1x = fmaf(qbias(wfx, kb, sel), sx, bx);
            const float t0y = fmaf(qbias(wny, kb, sel), sy, by), t1y = fmaf(qbias(wfy, kb, sel), sy, by);
            const float t0z = fmaf(qbias(wnz, kb, sel), sz, bz), t1z = fmaf(qbias(wfz, kb, sel), sz, bz);
            const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.f));
            const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            // branch-free bookkeeping on packed keys (t_near >= 0, so unsigned order = float order): keep the smallest
            // key, push the other one. The 3 cleared mantissa bits make t_near at most 7 ulp smaller: conservative.
            const uint32_t key = tn <= tf ? ((__float_as_uint(tn) & km) | (uint32_t) c) : 0xffffffffu;
            const uint32_t hi = max(best, key);
            best = min(best, key);
            // predicated store: an unconditional one costs no instruction less and was measured to fill L1 with
            // dead stack sectors (78 G local-store sectors per launch, local loads hitting L1 only 45 %)
            if (hi != 0xffffffffu) { st[sp] = make_uint2(xh, hi); ++sp; }
        }
        xh |= 0x80000000u;
    }
    // nearest of the two rounds continues, the other one is pushed
    const bool first = best0 <= best;
    const uint32_t nearest = first ? best0 : best, other = first ? best : best0;
    if (other != 0xffffffffu) { st[sp] = make_uint2(first ? xh : (xh & 0x7fffffffu), other); ++sp; }
    T.cur = nearest != 0xffffffffu ? trav_child_ref(n1, (nearest & 3u) | (first ? 0u : 4u)) : trav_pop(bvh, sp, st, tmax);
    T.sp = sp;
}

// One leaf (up to 3 triangles, Moeller-Trumbore as mesh.h:432-452), then pop. Requires T.cur to be a leaf.
// (Fetching the stack top and its node word ahead of the triangle tests was measured slower: 706 vs 718 Mpaths/s.)
LRT_D void trav_leaf_step(const BvhView &bvh, float3 o, float3 d, Trav &T, uint2 *st, bool any_hit, unsigned &work) {
    const uint32_t first = T.cur & 0x0fffffffu, count = (T.cur >> 28) & 3u;
    int sp = T.sp; float tmax = T.tmax;
    work += count << 16;
    for (uint32_t k = 0; k < count; ++k) {
        const float4 *tp = bvh.tris + 3 * (size_t) (first + k);
        const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
        const float3 p0 = make_float3(a.x, a.y, a.z), p1 = make_float3(a.w, b.x, b.y), p2 = make_float3(b.z, b.w, c.x);
        const float3 e1 = p1 - p0, e2 = p2 - p0;
        const float3 pvec = cross(d, e2);
        const float inv_det = 1.f / dot(e1, pvec);
        const float3 tvec = o - p0;
        const float u = dot(tvec, pvec) * inv_det;
        const float3 qvec = cross(tvec, e1);
        const float v = dot(d, qvec) * inv_det;
        const float t = dot(e2, qvec) * inv_det;
        const bool ok = u >= 0.f && u <= 1.f && v >= 0.f && u + v <= 1.f && t >= 0.f && t <= tmax;
        if (ok && (t < T.hit.t || (t == T.hit.t && (int) (first + k) < T.hit.tri))) {
            T.hit.t = t; T.hit.u = u; T.hit.v = v; T.hit.tri = (int) (first + k); tmax = t;
        }
    }
    T.cur = (any_hit && T.hit.tri >= 0) ? kTravDone : trav_pop(bvh, sp, st, tmax);
    T.sp = sp; T.tmax = tmax;
}

// A burst of traversal for a whole warp (all 32 lanes must call it). Every iteration is warp-uniform: EITHER all lanes
// standing on an inner node process that one node, OR (once `leaf_min` lanes wait on a leaf, or no lane has a node left)
// all lanes standing on a leaf test its triangles — node code and triangle code never run interleaved at low width.
// The burst ends when the number of traversing lanes has dropped by `exit_drop` 32nds of what it was at entry (at
// least by one): finished lanes go and shade / fetch new rays instead of idling until the slowest ray is done.
LRT_D void bvh_burst(const BvhView &bvh, float3 o, float3 d, Trav &T, uint2 *st, bool live, bool any_hit,
                     int exit_drop, int leaf_min, unsigned &work) {
    const float3 idir = trav_idir(d);
    const int a0 = __popc(__ballot_sync(0xffffffffu, live && T.cur != kTravDone));
    if (a0 == 0) return;
    const int floor_active = a0 - max(1, (a0 * exit_drop) >> 5);
    while (true) {
        const bool act = live && T.cur != kTravDone;
        const bool at_leaf = act && (T.cur & 0x80000000u) != 0u;
        const unsigned lm = __ballot_sync(0xffffffffu, at_leaf), nm = __ballot_sync(0xffffffffu, act && !at_leaf);
        if (__popc(lm | nm) <= floor_active) break;
        if (nm == 0u || __popc(lm) >= leaf_min) {
            if (at_leaf) trav_leaf_step(bvh, o, d, T, st, any_hit, work);
        } else {
            if (act && !at_leaf) trav_node_step(bvh, o, idir, T, st, work);
        }
    }
}

// One ray, run to completion (unit-level Scene::ray_intersect / ray_test).
LRT_D RayHit bvh_intersect(const BvhView &bvh, float3 o, float3 d, float tmax, bool any_hit, unsigned &work) {
    uint2 st[kTravStack];
    Trav T; trav_init(T, tmax);
    const float3 idir = trav_idir(d);
    while (T.cur != kTravDone) {
        if (T.cur & 0x80000000u) trav_leaf_step(bvh, o, d, T, st, any_hit, work);
        else trav_node_step(bvh, o, idir, T, st, work);
    }
    return T.hit;
}

}  // namespace lrt
