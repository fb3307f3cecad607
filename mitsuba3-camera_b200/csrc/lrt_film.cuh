// lensrt — film: reconstruction filters and the splat footprint of ImageBlock::put.
// Replaces src/render/imageblock.cpp:174-532 (box fast path :210-233, non-coalesced :264-412, coalesced
// :418-531 incl. `normalize`), src/rfilters/gaussian.cpp:55-104 (the CPU/LLVM polynomial — the oracle is a
// CPU build), tent.cpp:53-55, box.cpp:41-43.
#pragma once
#include "lrt_math.cuh"

namespace lrt {

enum : int { FILTER_BOX = 0, FILTER_TENT = 1, FILTER_GAUSSIAN = 2 };
constexpr int kMaxTaps = 8;   // per axis (gaussian stddev .5 -> 5 coalesced / 4 non-coalesced)

struct FilmParams {
    int size_x, size_y, off_x, off_y, border, channels;
    int kind, coalesce, normalize;
    int n_half, count;          // coalesced window: count = 2*n_half+1
    int count_max;              // non-coalesced: ceil(2*radius)
    float radius, inv_radius;
    float coeff[10];
    // scalar-variant splat (imageblock.cpp:264-379 with JIT = false): 31-tap table of ReconstructionFilter::
    // init_discretization (src/core/rfilter.cpp:9-24); `discretized` switches the non-coalesced footprint to it
    int discretized; float disc_scale; float disc[32];
};

LRT_HD float filter_eval(const FilmParams &F, float x) {
    if (F.kind == FILTER_GAUSSIAN) {
        // estrin_impl over 10 coefficients in x^2 (drjit/math.h:27-41)
        // (strict two-rounding fmadd: the float32 polynomial has ~1e-5 of cancellation noise near the filter
        //  edge, so the evaluation order is kept identical to the pinned CPU build)
        using strict::fmadd; using strict::sqr;
        const float *c = F.coeff;
        float x1 = sqr(x), x2 = sqr(x1), x4 = sqr(x2), x8 = sqr(x4);
        float a0 = fmadd(x1, c[1], c[0]), a1 = fmadd(x1, c[3], c[2]), a2 = fmadd(x1, c[5], c[4]),
              a3 = fmadd(x1, c[7], c[6]), a4 = fmadd(x1, c[9], c[8]);
        float b0 = fmadd(x2, a1, a0), b1 = fmadd(x2, a3, a2);
        float c0 = fmadd(x4, b1, b0);
        return fmaxf(fmadd(x8, a4, c0), 0.f);
    }
    if (F.kind == FILTER_TENT) return fmaxf(0.f, 1.f - fabsf(x * F.inv_radius));
    return (x >= -.5f && x < .5f) ? 1.f : 0.f;
}

// ReconstructionFilter::eval_discretized (include/mitsuba/core/rfilter.h:66-76)
LRT_HD float filter_eval_discretized(const FilmParams &F, float x) {
    const unsigned index = (unsigned) fabsf(x * F.disc_scale);
    return F.disc[index < 31u ? index : 31u];
}

// Footprint of one sample: window origin (x0,y0) in tensor coordinates, extent, separable weights.
struct Footprint {
    int x0, y0, nx, ny;         // taps ix in [0,nx) land on pixel x0+ix (still to be clipped to the tensor)
    float wx[kMaxTaps], wy[kMaxTaps];
};

template <bool kDisc = false> LRT_HD bool film_footprint(const FilmParams &F, float px, float py, Footprint &fp) {
    int sx = F.size_x + 2 * F.border, sy = F.size_y + 2 * F.border;
    if (F.kind == FILTER_BOX) {
        int ix = (int) floorf(px) - F.off_x, iy = (int) floorf(py) - F.off_y;
        if (ix < 0 || iy < 0 || ix >= F.size_x || iy >= F.size_y) return false;
        fp.x0 = ix; fp.y0 = iy; fp.nx = fp.ny = 1; fp.wx[0] = fp.wy[0] = 1.f;
        return true;
    }
    if (!F.coalesce) {
        float pfx = px + ((float) (F.border - F.off_x) - .5f), pfy = py + ((float) (F.border - F.off_y) - .5f);
        float p0x = pfx - F.radius, p0y = pfy - F.radius;
        int x0 = max((int) ceilf(p0x), 0), y0 = max((int) ceilf(p0y), 0);
        int x1 = min((int) floorf(pfx + F.radius), sx - 1), y1 = min((int) floorf(pfy + F.radius), sy - 1);
        if (x0 > x1 || y0 > y1) return false;
        fp.x0 = x0; fp.y0 = y0; fp.nx = min(x1 - x0 + 1, F.count_max); fp.ny = min(y1 - y0 + 1, F.count_max);
        float rx = (float) x0 - pfx, ry = (float) y0 - pfy;
        if (kDisc)   // (compile-time: the render megakernel never carries the table lookups)
            for (int i = 0; i < F.count_max; ++i) { fp.wx[i] = filter_eval_discretized(F, rx); fp.wy[i] = filter_eval_discretized(F, ry); rx += 1.f; ry += 1.f; }
        else
            for (int i = 0; i < F.count_max; ++i) { fp.wx[i] = filter_eval(F, rx); fp.wy[i] = filter_eval(F, ry); rx += 1.f; ry += 1.f; }
        if (F.normalize) {
            float a = 0.f, b = 0.f, r2x = ceilf(p0x) - pfx, r2y = ceilf(p0y) - pfy;
            for (int i = 0; i < F.count_max; ++i) {     // (scalar variants normalise with the discretised weights, imageblock.cpp:322-331)
                a += kDisc ? filter_eval_discretized(F, r2x) : filter_eval(F, r2x); b += kDisc ? filter_eval_discretized(F, r2y) : filter_eval(F, r2y);
                r2x += 1.f; r2y += 1.f;
            }
            float f = a * b; f = f != 0.f ? 1.f / f : 0.f;
            for (int i = 0; i < F.count_max; ++i) fp.wx[i] *= f;
        }
        return true;
    }
    int pix = (int) floorf(px) - F.n_half, piy = (int) floorf(py) - F.n_half;
    fp.x0 = pix + (F.border - F.off_x); fp.y0 = piy + (F.border - F.off_y); fp.nx = fp.ny = F.count;
    float rx = ((float) pix + .5f) - px, ry = ((float) piy + .5f) - py;
    for (int i = 0; i < F.count; ++i) { fp.wx[i] = filter_eval(F, rx); fp.wy[i] = filter_eval(F, ry); rx += 1.f; ry += 1.f; }
    if (F.normalize) {
        float a = 0.f, b = 0.f;
        for (int i = 0; i < F.count; ++i) { a += fp.wx[i]; b += fp.wy[i]; }
        float f = a * b; f = f != 0.f ? 1.f / f : 0.f;
        for (int i = 0; i < F.count; ++i) fp.wx[i] *= f;
    }
    return true;
}

}  // namespace lrt
