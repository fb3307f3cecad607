// lensrt — sample generator: TEA-scrambled PCG32 streams, one per Monte Carlo sample.
// Replaces: include/mitsuba/core/random.h:76-90 (sample_tea_32), ext/drjit/include/drjit/random.h:43-137
// (dr::PCG32) and the JIT branch of PCG32Sampler::seed (src/render/sampler.cpp:115-134).
// Integer-only, bit-exact by construction; the stream of sample `idx` depends only on (seed, idx), so
// any sharding of samples over lanes, CTAs or GPUs reproduces the same numbers.
#pragma once
#include "lrt_math.cuh"

namespace lrt {

LRT_HD void tea32(uint32_t v0, uint32_t v1, uint32_t &o0, uint32_t &o1, int rounds = 4) {
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}

struct Pcg32 {
    uint64_t state, inc;

    LRT_HD uint32_t next_u32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xs >> rot) | (xs << ((0u - rot) & 31u));
    }
    // [0,1): ((u >> 9) | 0x3f800000) - 1   (drjit/random.h:130-132)
    LRT_HD float next_f32() { return strict::sub(u2f((next_u32() >> 9) | 0x3f800000u), 1.f); }

    LRT_HD void seed(uint64_t initstate, uint64_t initseq) {
        state = 0;
        inc = (initseq << 1) | 1u;
        next_u32();
        state += initstate;
        next_u32();
    }
    // sampler->seed(seed, wavefront) lane `idx` (sampler.cpp:121-130)
    LRT_HD void seed_sample(uint32_t seed_value, uint32_t idx) {
        uint32_t v0, v1;
        tea32(seed_value, idx, v0, v1);
        seed((uint64_t) v0, (uint64_t) v1);
    }
};

}  // namespace lrt
