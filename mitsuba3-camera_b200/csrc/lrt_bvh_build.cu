// lensrt — host BVH builder: binned-SAH binary tree -> greedy collapse to 8-wide nodes -> quantised boxes.
// See lrt_bvh.h for the layout and the reference interfaces this replaces.
#include <vector_types.h>
#include "lrt_bvh.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>

namespace lrt {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<float>::infinity(); hi[a] = -lo[a]; } }
    void grow(const float *p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    void grow(const Box &b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    float area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return (dx < 0) ? 0.f : 2.f * (dx * dy + dy * dz + dz * dx);
    }
};

struct BinNode { Box box; int left = -1, right = -1, first = 0, count = 0; };

struct Builder {
    const float *verts; const int32_t *idx; int64_t nt;
    std::vector<Box> tbox; std::vector<float> cent;   // per triangle
    std::vector<int> order;
    std::vector<BinNode> bn;
    static constexpr int kBins = 16;
    int kLeaf = 3;

    int build(int first, int count, int depth, int &max_depth) {
        BinNode nd; nd.box.reset(); nd.first = first; nd.count = count;
        Box cb; cb.reset();
        for (int i = first; i < first + count; ++i) { nd.box.grow(tbox[order[i]]); cb.grow(&cent[3 * (size_t) order[i]]); }
        int id = (int) bn.size(); bn.push_back(nd);
        max_depth = std::max(max_depth, depth);
        if (count <= kLeaf) return id;
        // binned SAH over the longest centroid axis first, then the others
        float best_cost = std::numeric_limits<float>::infinity(); int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; ++axis) {
            float lo = cb.lo[axis], ext = cb.hi[axis] - lo;
            if (!(ext > 0.f)) continue;
            Box bb[kBins]; int bc[kBins];
            for (int b = 0; b < kBins; ++b) { bb[b].reset(); bc[b] = 0; }
            float k = kBins * (1.f - 1e-6f) / ext;
            for (int i = first; i < first + count; ++i) {
                int t = order[i]; int b = std::min(kBins - 1, std::max(0, (int) ((cent[3 * (size_t) t + axis] - lo) * k)));
                bb[b].grow(tbox[t]); bc[b]++;
            }
            float ra[kBins]; Box acc; acc.reset(); int cnt = 0;
            int rc[kBins];
            for (int b = kBins - 1; b > 0; --b) { acc.grow(bb[b]); cnt += bc[b]; ra[b] = acc.area(); rc[b] = cnt; }
            acc.reset(); cnt = 0;
            for (int b = 0; b < kBins - 1; ++b) {
                acc.grow(bb[b]); cnt += bc[b];
                if (cnt == 0 || rc[b + 1] == 0) continue;
                float cost = acc.area() * cnt + ra[b + 1] * rc[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        int mid;
        if (best_axis < 0) {   // all centroids coincide: split in the middle
            mid = first + count / 2;
        } else {
            float lo = cb.lo[best_axis], ext = cb.hi[best_axis] - lo, k = kBins * (1.f - 1e-6f) / ext;
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](int t) {
                int b = std::min(kBins - 1, std::max(0, (int) ((cent[3 * (size_t) t + best_axis] - lo) * k)));
                return b <= best_bin;
            });
            mid = (int) (it - order.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        int l = build(first, mid - first, depth + 1, max_depth);
        int r = build(mid, first + count - mid, depth + 1, max_depth);
        bn[id].left = l; bn[id].right = r; bn[id].count = 0;
        return id;
    }
};

// Optimal collapse of the binary tree into 8-wide nodes (the dynamic programme of Ylitie, Karras & Laine, "Efficient
// Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017, section 3.1): cost[n][i] = cheapest way to
// represent the subtree of binary node n as a forest of at most i+1 wide-node children (i = 0..6 -> 1..7 slots).
//   C(n, 1) = min(C_leaf(n), C_internal(n)),  C_internal(n) = A(n) c_node + min_k C(l, k) + C(r, 8 - k)
//   C(n, i) = min(C(n, i - 1), min_k C(l, k) + C(r, i - k))
// The greedy "open the largest child" collapse left 60 % of the nodes of a 1M-triangle height field with only 2-3 children
// (subtrees of 2-3 leaves that never got opened because their parent's slots had run out); the optimum fills the bottom.
struct Collapse {
    const std::vector<BinNode> &bn;
    static constexpr float kNodeCost = 1.f, kTriCost = 0.3f;
    std::vector<float> cost;        // [n][7]
    std::vector<uint8_t> dec;       // [n][7]: 0 = leaf (i = 1 only), 255 = internal node (i = 1 only), 254 = same as i - 1, else k = slots of the left child
    std::vector<uint8_t> split8;    // [n]: left-child slots of the best 8-slot distribution (the node itself as a wide node)
    explicit Collapse(const std::vector<BinNode> &b) : bn(b), cost(b.size() * 7), dec(b.size() * 7), split8(b.size(), 0) {
        for (int n = (int) bn.size() - 1; n >= 0; --n) solve(n);      // children have larger indices than their parent (pre-order)
    }
    float C(int n, int i) const { return cost[(size_t) n * 7 + (i - 1)]; }
    void solve(int n) {
        const BinNode &b = bn[n];
        float *c = &cost[(size_t) n * 7]; uint8_t *d = &dec[(size_t) n * 7];
        const float A = b.box.area();
        if (b.left < 0) {                 // binary leaf (<= 3 triangles)
            for (int i = 0; i < 7; ++i) { c[i] = A * kTriCost * (float) b.count; d[i] = i == 0 ? 0 : 254; }
            return;
        }
        const int l = b.left, r = b.right;
        float best8 = std::numeric_limits<float>::infinity(); int k8 = 1;
        for (int k = 1; k <= 7; ++k) { const float v = C(l, k) + C(r, 8 - k); if (v < best8) { best8 = v; k8 = k; } }
        split8[n] = (uint8_t) k8;
        c[0] = A * kNodeCost + best8; d[0] = 255;
        for (int i = 2; i <= 7; ++i) {
            float best = c[i - 2]; uint8_t how = 254;
            for (int k = 1; k < i; ++k) { const float v = C(l, k) + C(r, i - k); if (v < best) { best = v; how = (uint8_t) k; } }
            c[i - 1] = best; d[i - 1] = how;
        }
    }
    // the (binary) nodes that fill `i` slots under n
    void collect(int n, int i, int *out, int &cnt) const {
        while (i > 1 && dec[(size_t) n * 7 + (i - 1)] == 254) --i;
        if (i == 1) { out[cnt++] = n; return; }
        const int k = dec[(size_t) n * 7 + (i - 1)];
        collect(bn[n].left, k, out, cnt); collect(bn[n].right, i - k, out, cnt);
    }
    int children(int n, int *out) const {     // children of the wide node rooted at the inner binary node n
        int cnt = 0;
        collect(bn[n].left, split8[n], out, cnt); collect(bn[n].right, 8 - split8[n], out, cnt);
        return cnt;
    }
};

uint8_t exponent_for(float extent) {
    // smallest e with 253 * 2^(e-127) >= extent (one quantisation step is reserved on each side, see below)
    if (!(extent > 0.f)) return 1;
    int e; std::frexp(extent / 253.f, &e);   // extent/253 = m * 2^e, m in [.5,1)  =>  2^e >= extent/253
    int biased = e + 127;
    return (uint8_t) std::min(254, std::max(1, biased));
}

}  // namespace

void build_bvh8(const float *verts, const int32_t *idx, const int32_t *tri_shape, int64_t nt,
                std::vector<BvhNode> &nodes, std::vector<BvhTri> &tris, BvhBuildStats *stats) {
    nodes.clear(); tris.clear();
    Builder B; B.verts = verts; B.idx = idx; B.nt = nt;
    if (const char *e = getenv("LRT_BVH_LEAF")) B.kLeaf = std::min(3, std::max(1, atoi(e)));
    B.tbox.resize(nt); B.cent.resize(3 * (size_t) nt); B.order.resize(nt);
    for (int64_t t = 0; t < nt; ++t) {
        Box b; b.reset();
        for (int k = 0; k < 3; ++k) b.grow(verts + 3 * (size_t) idx[3 * t + k]);
        B.tbox[t] = b; B.order[t] = (int) t;
        for (int a = 0; a < 3; ++a) B.cent[3 * t + a] = 0.5f * (b.lo[a] + b.hi[a]);
    }
    int max_depth = 0;
    if (nt == 0) {   // empty scene: a root whose children are all empty
        BvhNode n; memset(&n, 0, sizeof n); n.ex = n.ey = n.ez = 1;
        memset(n.qlo_x, 255, 24);   // qlo_x, qlo_y, qlo_z: inverted boxes (lo 255, hi 0) never pass the slab test
        nodes.push_back(n);
        if (stats) { *stats = BvhBuildStats(); stats->level_begin = { 0, 1 }; }
        return;
    }
    B.bn.reserve(2 * (size_t) nt);
    B.build(0, (int) nt, 0, max_depth);

    const char *cm = getenv("LRT_BVH_COLLAPSE");
    const bool optimal = !(cm && cm[0] == 'g');          // LRT_BVH_COLLAPSE=greedy restores the round-1 collapse
    Collapse *dp = optimal ? new Collapse(B.bn) : nullptr;
    // collapse: breadth-first emission so that the inner children of a node are contiguous
    struct Pending { int bin; };   // binary subtree root that becomes one wide node
    std::vector<int> queue; queue.push_back(0);
    nodes.resize(1);
    size_t head = 0; int leaves = 0;
    std::vector<int> level_begin; level_begin.push_back(0);
    size_t level_end = 1;          // first queue index of the next level
    while (head < queue.size()) {
        if (head == level_end) { level_begin.push_back((int) head); level_end = queue.size(); }
        int bin_root = queue[head]; size_t self = head; head++;
        // gather up to 8 children by repeatedly opening the inner child with the largest area
        int ch[8]; int nch = 0;
        const BinNode &root = B.bn[bin_root];
        if (root.left < 0) { ch[nch++] = bin_root; }
        else if (dp) nch = dp->children(bin_root, ch);
        else { ch[nch++] = root.left; ch[nch++] = root.right; }
        while (!dp && nch < 8) {
            int best = -1; float ba = -1.f;
            for (int i = 0; i < nch; ++i) if (B.bn[ch[i]].left >= 0) { float a = B.bn[ch[i]].box.area(); if (a > ba) { ba = a; best = i; } }
            if (best < 0) break;
            int c = ch[best]; ch[best] = B.bn[c].left; ch[nch++] = B.bn[c].right;
        }
        BvhNode n; memset(&n, 0, sizeof n);
        memset(n.qlo_x, 255, 24);   // unused child slots keep an inverted box (lo 255, hi 0): the traversal relies on it
        const Box &nb = root.box;
        n.ex = exponent_for(nb.hi[0] - nb.lo[0]); n.ey = exponent_for(nb.hi[1] - nb.lo[1]); n.ez = exponent_for(nb.hi[2] - nb.lo[2]);
        n.child_base = (uint32_t) queue.size(); n.tri_base = (uint32_t) tris.size();
        float sc[3] = { std::ldexp(1.f, n.ex - 127), std::ldexp(1.f, n.ey - 127), std::ldexp(1.f, n.ez - 127) };
        // the grid starts one step below the box so that every child bound keeps a full step of slack on both
        // sides (q_lo = floor(x) - 1 >= 0, q_hi = ceil(x) + 1 <= 255): this absorbs the float rounding of the
        // quantisation and of the biased-FMA decode in the traversal (lrt_trace.cuh)
        n.ox = nb.lo[0] - sc[0]; n.oy = nb.lo[1] - sc[1]; n.oz = nb.lo[2] - sc[2];
        int inner_k = 0, tri_off = 0;
        for (int i = 0; i < nch; ++i) {
            const BinNode &c = B.bn[ch[i]];
            uint8_t *ql[3] = { n.qlo_x, n.qlo_y, n.qlo_z }, *qh[3] = { n.qhi_x, n.qhi_y, n.qhi_z };
            const float org[3] = { n.ox, n.oy, n.oz };
            for (int a = 0; a < 3; ++a) {
                float lo = std::floor((c.box.lo[a] - org[a]) / sc[a]), hi = std::ceil((c.box.hi[a] - org[a]) / sc[a]);
                ql[a][i] = (uint8_t) std::min(255.f, std::max(0.f, lo - 1.f));
                qh[a][i] = (uint8_t) std::min(255.f, std::max(0.f, hi + 1.f));
            }
            // (optimal collapse: a child is an inner wide node iff the programme decided so; a binary leaf is always a leaf)
            if (c.left >= 0) { n.meta[i] = (uint8_t) (0x80 | inner_k); inner_k++; queue.push_back(ch[i]); }
            else {
                n.meta[i] = (uint8_t) ((c.count << 5) | tri_off); tri_off += c.count; leaves++;
                for (int k = c.first; k < c.first + c.count; ++k) {
                    int t = B.order[k]; BvhTri tr; memset(&tr, 0, sizeof tr);
                    for (int a = 0; a < 3; ++a) {
                        tr.p0[a] = verts[3 * (size_t) idx[3 * (size_t) t] + a];
                        tr.p1[a] = verts[3 * (size_t) idx[3 * (size_t) t + 1] + a];
                        tr.p2[a] = verts[3 * (size_t) idx[3 * (size_t) t + 2] + a];
                    }
                    tr.shape = tri_shape ? tri_shape[t] : 0; tr.prim = t;
                    tris.push_back(tr);
                }
            }
        }
        if (nodes.size() < queue.size()) nodes.resize(queue.size());
        nodes[self] = n;
    }
    delete dp;
    level_begin.push_back((int) nodes.size());
    if (stats) { stats->nodes = (int) nodes.size(); stats->leaves = leaves; stats->max_depth = max_depth; stats->level_begin = level_begin; }
}

}  // namespace lrt
