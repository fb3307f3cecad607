// lensrt — scene handle: device-resident BVH + shape / BSDF / emitter tables
#pragma once
#include "lrt_internal.h"
#include "lrt_shade.cuh"
#include <vector>

namespace lrt { constexpr int kCounterRing = 64; }

struct lrt_scene {
    lrt::SceneView view;
    void *d_nodes = nullptr, *d_tris = nullptr, *d_shape_bsdf = nullptr, *d_shape_emitter = nullptr,
         *d_bsdfs = nullptr, *d_emitters = nullptr, *d_tri_attr = nullptr, *d_tex_data = nullptr, *d_textures = nullptr, *d_em_tris = nullptr, *d_spheres = nullptr;
    // refit support (lrt_refit.cu): vertex indices of every reordered triangle, per-node float boxes, BFS level ranges
    void *d_tri_vidx = nullptr, *d_node_box = nullptr; std::vector<int> level_begin; int64_t n_vertices = 0;
    unsigned long long *d_counter = nullptr;   // ring of work counters [0..kCounterRing), then 8 statistics words
    mutable std::atomic<unsigned> launch_seq { 0 };
    int64_t n_nodes = 0, n_tris = 0; int n_shapes = 0, n_bsdfs = 0, n_emitters = 0, n_textures = 0, bvh_depth = 0;
    bool has_directional = false;         // some emitter is a `directionalarea` (row[14] = 1): the particle tracer only
    double build_seconds = 0;
};
