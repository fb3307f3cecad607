// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). Film step restatement:
// ImageBlock::put (src/render/imageblock.cpp:174-532), reconstruction filters
// (src/rfilters/gaussian.cpp:48-104, tent.cpp:47-55, box), HDRFilm::develop (src/films/hdrfilm.cpp:304-418).
// JIT conventions (continuous filter evaluation; coalesced window for spp >= 4). Accumulation is done in
// float64 so that the checker is order-independent; the product accumulates float32 atomics.
#include "oracle.h"
#include "oracle_film.h"
#include <cstring>
#include <algorithm>

namespace orc {

static int ceil2int(float x) { return (int) std::ceil(x); }
static int floor2int(float x) { return (int) std::floor(x); }

// imageblock.cpp:174-532; tensor is (size+2*border)^2 x channels, float64 accumulators
void block_put(const Filter &flt, int size_x, int size_y, int off_x, int off_y, int border, int channels,
               bool coalesce, bool normalize, bool discretized, float pos_x, float pos_y, const float *values, double *tensor) {
    if (flt.kind == 0) {   // :210-233 (m_rfilter == nullptr for the box filter)
        int px = floor2int(pos_x) - off_x, py = floor2int(pos_y) - off_y;
        if (px < 0 || py < 0 || px >= size_x || py >= size_y) return;
        size_t index = ((size_t) py * size_x + px) * channels;
        for (int k = 0; k < channels; ++k) tensor[index + k] += values[k];
        return;
    }
    float radius = flt.radius;
    int sx = size_x + 2 * border, sy = size_y + 2 * border;
    if (!coalesce || discretized) {       // :264-412 (JIT: continuous filter, conservative count; scalar: discretised filter)
        float pfx = pos_x + ((float) (border - off_x) - .5f), pfy = pos_y + ((float) (border - off_y) - .5f);
        float p0x = pfx - radius, p0y = pfy - radius, p1x = pfx + radius, p1y = pfy + radius;
        int x0 = std::max(ceil2int(p0x), 0), y0 = std::max(ceil2int(p0y), 0);
        int x1 = std::min(floor2int(p1x), sx - 1), y1 = std::min(floor2int(p1y), sy - 1);
        if (x0 > x1 || y0 > y1) return;
        int count_max = ceil2int(2.f * radius);
        float wx[16], wy[16];
        float rx = (float) x0 - pfx, ry = (float) y0 - pfy;
        auto ev = [&](float x) { return discretized ? flt.eval_discretized(x) : flt.eval(x); };
        for (int i = 0; i < count_max; ++i) { wx[i] = ev(rx); rx += 1.f; wy[i] = ev(ry); ry += 1.f; }
        if (normalize) {   // :318-342
            float sxw = 0.f, syw = 0.f;
            float r2x = std::ceil(p0x) - pfx, r2y = std::ceil(p0y) - pfy;
            for (int i = 0; i < count_max; ++i) { sxw += ev(r2x); syw += ev(r2y); r2x += 1.f; r2y += 1.f; }
            float factor = sxw * syw;
            factor = factor != 0.f ? rcp(factor) : 0.f;
            for (int i = 0; i < count_max; ++i) wx[i] *= factor;
        }
        for (int y = 0; y < count_max; ++y) {
            if (y0 + y > y1) continue;
            for (int x = 0; x < count_max; ++x) {
                if (x0 + x > x1) continue;
                float w = wx[x] * wy[y];
                size_t index = ((size_t) (y0 + y) * sx + (x0 + x)) * channels;
                for (int k = 0; k < channels; ++k) tensor[index + k] += (double) (values[k] * w);
            }
        }
        return;
    }
    // coalesced (:418-531): fixed (2n+1)^2 window anchored at floor(pos) - n, weights at pixel centres
    int n = ceil2int(radius - .5f), count = 2 * n + 1;
    int pix = floor2int(pos_x) - n, piy = floor2int(pos_y) - n;
    int lx = pix + (border - off_x), ly = piy + (border - off_y);
    float rx = ((float) pix + .5f) - pos_x, ry = ((float) piy + .5f) - pos_y;
    float wx[16], wy[16];
    for (int i = 0; i < count; ++i) { wx[i] = flt.eval(rx); wy[i] = flt.eval(ry); rx += 1.f; ry += 1.f; }
    if (normalize) {
        float a = 0.f, b = 0.f;
        for (int i = 0; i < count; ++i) { a += wx[i]; b += wy[i]; }
        float factor = a * b;
        factor = factor != 0.f ? rcp(factor) : 0.f;
        for (int i = 0; i < count; ++i) wx[i] *= factor;
    }
    for (int ys = 0; ys < count; ++ys) {
        unsigned y = (unsigned) (ly + ys);
        if (y >= (unsigned) sy) continue;
        for (int xs = 0; xs < count; ++xs) {
            unsigned x = (unsigned) (lx + xs);
            if (x >= (unsigned) sx) continue;
            float w = wy[ys] * wx[xs];
            size_t index = ((size_t) y * sx + x) * channels;
            for (int k = 0; k < channels; ++k) tensor[index + k] += (double) (values[k] * w);
        }
    }
}

}  // namespace orc

using namespace orc;
extern "C" {
void orc_filter_eval(int kind, float param, int n, const float *x, float *out) {
    Filter f; f.init(kind, param);
    for (int i = 0; i < n; ++i) out[i] = f.eval(x[i]);
}
float orc_filter_radius(int kind, float param) { Filter f; f.init(kind, param); return f.radius; }
// n samples: pos[n,2], values[n,channels] -> tensor64 [(size_y+2b),(size_x+2b),channels] (accumulated into)
void orc_block_put(int kind, float param, int size_x, int size_y, int off_x, int off_y, int border, int channels,
                   int coalesce, int normalize, int64_t n, const float *pos, const float *values, double *tensor64) {
    Filter f; f.init(kind, param);
    for (int64_t i = 0; i < n; ++i)
        block_put(f, size_x, size_y, off_x, off_y, border, channels, coalesce != 0, normalize != 0, false,
                  pos[2 * i], pos[2 * i + 1], values + (size_t) i * channels, tensor64);
}
// hdrfilm.cpp:390-392 develop for pixel_format rgb: storage [H,W,4] (R,G,B,W) -> [H,W,3]
void orc_develop_rgb(int64_t npix, const double *storage4, float *rgb) {
    for (int64_t i = 0; i < npix; ++i) {
        float w = (float) storage4[4 * i + 3];
        float d = w == 0.f ? 1.f : w;
        for (int k = 0; k < 3; ++k) rgb[3 * i + k] = (float) storage4[4 * i + k] / d;
    }
}
}
