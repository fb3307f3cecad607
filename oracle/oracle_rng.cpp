// oracle/ — TEST INFRASTRUCTURE ONLY (see oracle.h). RNG restatement.
#include "oracle.h"

namespace orc {

// include/mitsuba/core/random.h:76-90 — Tiny Encryption Algorithm, `rounds` Feistel rounds
void tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t &o0, uint32_t &o1) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}

// ext/drjit/include/drjit/random.h:55-63 (size == 1: arange(1) == 0)
void PCG32::seed(uint64_t initstate, uint64_t initseq) {
    state = 0;
    inc = (initseq << 1) | 1u;
    next_uint32();
    state += initstate;
    next_uint32();
}

// ext/drjit/include/drjit/random.h:66-75
uint32_t PCG32::next_uint32() {
    uint64_t oldstate = state;
    state = oldstate * 0x5851f42d4c957f2dULL + inc;
    uint32_t xorshifted = (uint32_t) (((oldstate >> 18u) ^ oldstate) >> 27u);
    uint32_t rot = (uint32_t) (oldstate >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((-(int32_t) rot) & 31));
}

// ext/drjit/include/drjit/random.h:130-132
float PCG32::next_float32() {
    union { uint32_t u; float f; } x;
    x.u = (next_uint32() >> 9) | 0x3f800000u;
    return x.f - 1.f;
}

// src/render/sampler.cpp:115-134, JIT branch: (v0, v1) = tea(seed_value, idx); rng.seed(1, v0, v1)
void sampler_seed_jit(PCG32 &rng, uint32_t seed, uint32_t idx) {
    uint32_t v0, v1;
    tea32(seed, idx, 4, v0, v1);
    rng.seed((uint64_t) v0, (uint64_t) v1);
}

}  // namespace orc

extern "C" {
void orc_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) {
    orc::tea32(v0, v1, rounds, out[0], out[1]);
}
// include/mitsuba/core/random.h:136-140 sample_tea_float32 (uses the SECOND word)
float orc_tea_float32(uint32_t v0, uint32_t v1, int rounds) {
    uint32_t a, b; orc::tea32(v0, v1, rounds, a, b);
    union { uint32_t u; float f; } x; x.u = (b >> 9) | 0x3f800000u;
    return x.f - 1.f;
}
void orc_pcg32_u32(uint64_t initstate, uint64_t initseq, int n, uint32_t *out) {
    orc::PCG32 r; r.seed(initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = r.next_uint32();
}
void orc_pcg32_f32(uint64_t initstate, uint64_t initseq, int n, float *out) {
    orc::PCG32 r; r.seed(initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = r.next_float32();
}
// first `ndraw` floats of each of the lanes [idx0, idx0+n) of a JIT-seeded sampler
void orc_sampler_jit_f32(uint32_t seed, uint32_t idx0, int n, int ndraw, float *out) {
    for (int i = 0; i < n; ++i) {
        orc::PCG32 r; orc::sampler_seed_jit(r, seed, idx0 + (uint32_t) i);
        for (int k = 0; k < ndraw; ++k) out[(size_t) i * ndraw + k] = r.next_float32();
    }
}
}
