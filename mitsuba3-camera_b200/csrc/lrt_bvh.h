// lensrt — 8-wide BVH with quantised child boxes (layout shared by the host builder and the device traversal).
// Stands in for the reference's three acceleration back ends (native kd-tree src/render/kdtree.cpp +
// include/mitsuba/render/kdtree.h, Embree scene_embree.inl, OptiX scene_optix.inl:200-530) behind
// Scene::ray_intersect / ray_test (src/render/scene.cpp:181-210). None of them is reused.
//
// Node = 80 bytes = five 128-bit loads:
//   word 0   : origin.x, origin.y, origin.z (f32), ex | ey<<8 | ez<<16 | reserved<<24   (grid = origin + q * 2^(e-127))
//   word 1   : child_base (first inner child), tri_base (first leaf triangle), meta[0..3], meta[4..7]
//   words 2-4: qlo_x[8] qlo_y[8] | qlo_z[8] qhi_x[8] | qhi_y[8] qhi_z[8]   (one byte per child and bound)
// meta byte per child: 0 = empty; 0x80|k = inner child number k (node index child_base + k);
//                      (count << 5) | offset = leaf with `count` (1..3) triangles at tri_base + offset (< 32).
// Triangles = 48 bytes = three 128-bit loads: p0.xyz, p1.xyz, p2.xyz, shape id, original primitive id, pad.
#pragma once
#include <cstdint>
#include <vector>

namespace lrt {

struct alignas(16) BvhNode {
    float ox, oy, oz; uint8_t ex, ey, ez, pad;
    uint32_t child_base, tri_base; uint8_t meta[8];
    uint8_t qlo_x[8], qlo_y[8], qlo_z[8], qhi_x[8], qhi_y[8], qhi_z[8];
};
static_assert(sizeof(BvhNode) == 80, "BvhNode layout");

struct alignas(16) BvhTri {
    float p0[3], p1[3], p2[3]; int32_t shape, prim; float pad;
};
static_assert(sizeof(BvhTri) == 48, "BvhTri layout");

struct BvhBuildStats {
    int nodes = 0, leaves = 0, max_depth = 0; double sah = 0;
    // nodes are emitted breadth-first: level l = nodes [level_begin[l], level_begin[l+1]); every child index > its parent's
    // (what the device refit walks bottom-up, lrt_refit.cu)
    std::vector<int> level_begin;
};

// verts [nv,3], idx [nt,3], tri_shape [nt]; outputs the wide nodes and the reordered triangle array
void build_bvh8(const float *verts, const int32_t *idx, const int32_t *tri_shape, int64_t nt,
                std::vector<BvhNode> &nodes, std::vector<BvhTri> &tris, BvhBuildStats *stats);

// device-side view
struct BvhView {
    const uint4 *nodes;   // 5 x uint4 per node
    const float4 *tris;   // 3 x float4 per triangle
    // two constants the traversal wants in REGISTERS rather than as immediates (0x4B000000 and ~7, lrt_trace.cuh):
    // passed as data so that ptxas cannot fold them back into the instructions
    uint32_t kb, km;
};

}  // namespace lrt
