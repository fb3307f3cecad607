// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). Reconstruction filters shared by oracle_film.cpp and
// oracle_spot.cpp (src/rfilters/gaussian.cpp:48-104, tent.cpp:47-55, box.cpp).
#pragma once
#include "oracle.h"

namespace orc {

// drjit/math.h:27-41 estrin_impl, recursive pairing in x, x^2, x^4...
inline float estrin(float x, const float *coeff, int n) {
    float buf[16];
    for (int i = 0; i < n; ++i) buf[i] = coeff[i];
    while (n > 1) {
        int n_rec = (n - 1) / 2, n_fma = n / 2;
        float nxt[16];
        for (int i = 0; i < n_fma; ++i) nxt[i] = fmadd(x, buf[2 * i + 1], buf[2 * i]);
        if (n_rec == n_fma) nxt[n_rec] = buf[n - 1];
        n = n_rec + 1;
        for (int i = 0; i < n; ++i) buf[i] = nxt[i];
        x = sqr(x);
    }
    return buf[0];
}

struct Filter {
    int kind;         // 0 box, 1 tent, 2 gaussian
    float radius;
    float coeff[10];  // gaussian Remez coefficients (CPU/LLVM branch, gaussian.cpp:55-89)
    float inv_radius;
    float values[32], scale_factor;   // init_discretization (src/core/rfilter.cpp:9-24), MI_FILTER_RESOLUTION = 31
    void init(int k, float param) {
        kind = k; inv_radius = 0.f;
        if (k == 0) radius = 0.5f;
        if (k == 1) { radius = param > 0 ? param : 1.f; inv_radius = 1.f / radius; }
        if (k == 2) {
            float stddev = param > 0 ? param : 0.5f;
            radius = 4 * stddev;
            const double c[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                                   -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
            double scale = 1;
            for (int i = 0; i < 10; ++i) { coeff[i] = (float) (c[i] * scale); scale /= (double) stddev * (double) stddev; }
            coeff[0] -= estrin(sqr(radius), coeff, 10);
        }
        for (int i = 0; i < 31; ++i) values[i] = eval((radius * (float) i) / 31.f);
        values[31] = 0.f;
        scale_factor = 31.f / radius;
    }
    // include/mitsuba/core/rfilter.h:66-76 (scalar variants): index = min(uint(|x * scale|), 31)
    float eval_discretized(float x) const {
        uint32_t index = (uint32_t) std::fabs(x * scale_factor);
        return values[index < 31u ? index : 31u];
    }
    float eval(float x) const {
        if (kind == 1) return std::fmax(0.f, 1.f - std::fabs(x * inv_radius));                 // tent.cpp:53-55
        if (kind == 2) return std::fmax(estrin(sqr(x), coeff, 10), 0.f);                       // gaussian.cpp:94-96
        return (x >= -0.5f && x < 0.5f) ? 1.f : 0.f;                                            // box.cpp eval
    }
};


// imageblock.cpp:174-532; tensor is (size+2*border)^2 x channels, float64 accumulators. discretized = the scalar variants'
// path: non-coalesced window, exact tap counts, eval_discretized (imageblock.cpp:264-379 with JIT = false)
void block_put(const Filter &flt, int size_x, int size_y, int off_x, int off_y, int border, int channels,
               bool coalesce, bool normalize, bool discretized, float pos_x, float pos_y, const float *values, double *tensor);

}  // namespace orc
