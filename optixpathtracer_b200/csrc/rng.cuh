// rng.cuh — random-number policies of the path-tracing core.
//
//  * LcgRng<HOST_ORDER>: bit-compatible with the reference generator
//    (Renderer/OptiX/random.h:34-69: tea<N> hash, 24-bit LCG, rnd in [0,1)).
//    Two flavours reproduce the two compiler-dependent behaviours of the
//    reference source:
//      - order of the two draws in `glm::vec2(rnd(s), rnd(s))`
//        (random.h:79, PBRT/LambertDiffuse.h:40): nvcc = left-to-right,
//        g++ = right-to-left;
//      - `unsigned(float)` of a negative value in the layered BSDF's re-seed
//        (PBRT/GlossyDiffuse.h:215-217,417): PTX cvt.rzi.u32.f32 saturates to
//        0, x86-64 wraps modulo 2^32.
//  * PhiloxRng: production generator, Philox4x32-10, counter-based, so a
//    sample's stream depends only on (seed, pixel, sample index, bounce, stage)
//    and never on scheduling: renders are reproducible and the N-GPU split of
//    samples reproduces the 1-GPU image.
#pragma once

#include "vecmath.cuh"

namespace b2pt {

// random.h:34-48
template <unsigned N>
B2PT_HD uint32_t tea(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
#pragma unroll
    for (unsigned n = 0; n < N; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

// float -> unsigned the way the two compilers of the reference do it.
template <bool HOST_ORDER>
B2PT_HD uint32_t ref_f2u(float x) {
    if (HOST_ORDER) {
        // x86-64: cvttss2si to 64 bit, low half kept -> wraps for negatives.
        return (uint32_t)(long long)x;
    } else {
        // PTX cvt.rzi.u32.f32: NaN -> 0, negative -> 0, too large -> 0xffffffff.
        if (!(x > 0.0f)) return 0u;
        if (x >= 4294967296.0f) return 0xffffffffu;
        return (uint32_t)x;
    }
}

template <bool HOST_ORDER>
struct LcgRng {
    uint32_t s;

    B2PT_HD explicit LcgRng(uint32_t seed) : s(seed) {}

    // random.h:51-57, 66-69
    B2PT_HD float next() {
        s = 1664525u * s + 1013904223u;
        return (float)(s & 0x00FFFFFFu) / (float)0x01000000;
    }
    // the pair `glm::vec2(rnd(s), rnd(s))`: x = .x component, y = .y component
    B2PT_HD void next2(float& x, float& y) {
        if (HOST_ORDER) { y = next(); x = next(); }
        else            { x = next(); y = next(); }
    }
    // GlossyDiffuse.h:215-218 (f) — the side stream used only for the walk's
    // Russian roulette.
    B2PT_HD LcgRng fork4(float3 wo, float3 wi) const {
        uint32_t n = tea<16>(ref_f2u<HOST_ORDER>(wo.x * 1000), ref_f2u<HOST_ORDER>(wo.y * 1000));
        n = tea<16>(n, ref_f2u<HOST_ORDER>(wi.x * 1000));
        n = tea<16>(n, ref_f2u<HOST_ORDER>(wi.y * 1000));
        n = tea<16>(n, s);
        return LcgRng(n);
    }
    // GlossyDiffuse.h:417-418 (Sample_f)
    B2PT_HD LcgRng fork2(float3 wo) const {
        uint32_t n = tea<16>(ref_f2u<HOST_ORDER>(wo.x * 1000), ref_f2u<HOST_ORDER>(wo.y * 1000));
        n = tea<16>(n, s);
        return LcgRng(n);
    }
};

// Philox4x32-R (Salmon et al., SC'11); R = 10 is the generator's standard strength (7 is the paper's smallest
// Crush-resistant round count: the walk engine's A/B, fast_layered.cuh)
template <int ROUNDS>
B2PT_HD void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                        uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < ROUNDS; r++) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

B2PT_HD void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                           uint32_t k0, uint32_t k1, uint32_t out[4]) {
    philox4x32<10>(c0, c1, c2, c3, k0, k1, out);
}

// one 4-draw block; a real call on the device (ten rounds x every draw site is a lot of code)
B2PT_HD_CALL uint4 philox_block(uint32_t block, uint32_t c1, uint32_t c2, uint32_t k0, uint32_t k1) {
    uint32_t o[4];
    philox4x32_10(block, c1, c2, 0x62327074u /* "b2pt" */, k0, k1, o);
    return make_uint4(o[0], o[1], o[2], o[3]);
}
// the same block inlined at its one call site per walk step (the walk engine: one draw site, a small loop)
template <int ROUNDS>
B2PT_HD uint4 philox_block_inline(uint32_t block, uint32_t c1, uint32_t c2, uint32_t k0, uint32_t k1) {
    uint32_t o[4];
    philox4x32<ROUNDS>(block, c1, c2, 0x62327074u /* "b2pt" */, k0, k1, o);
    return make_uint4(o[0], o[1], o[2], o[3]);
}

struct PhiloxRng {
    uint32_t k0, k1;      // key: seed ^ pixel, seed_hi ^ sample index
    uint32_t c1, c2;      // counter words: (bounce | stage << 16), stream id
    uint32_t block;       // counter word 0: next 4-draw block
    uint32_t buf[4];
    uint32_t have;        // draws left in buf

    B2PT_HD PhiloxRng(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce_stage,
                      uint32_t stream = 0)
        : k0((uint32_t)seed ^ pixel), k1((uint32_t)(seed >> 32) ^ sample),
          c1(bounce_stage), c2(stream), block(0), have(0) {
        buf[0] = buf[1] = buf[2] = buf[3] = 0;
    }

    B2PT_HD uint32_t next_u32() {
        if (have == 0) {
            const uint4 o = philox_block(block++, c1, c2, k0, k1);
            buf[0] = o.x; buf[1] = o.y; buf[2] = o.z; buf[3] = o.w;
            have = 4;
        }
        // select without dynamic indexing (keeps buf in registers)
        const uint32_t i = 4 - have;
        have--;
        return i == 0 ? buf[0] : i == 1 ? buf[1] : i == 2 ? buf[2] : buf[3];
    }
    B2PT_HD float next() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }
    B2PT_HD void next2(float& x, float& y) { x = next(); y = next(); }
    // side streams of the layered BSDF: same key, distinct stream word
    B2PT_HD PhiloxRng fork4(float3, float3) const { return fork_(c2 * 2 + 1); }
    B2PT_HD PhiloxRng fork2(float3) const { return fork_(c2 * 2 + 2); }

  private:
    B2PT_HD PhiloxRng fork_(uint32_t stream) const {
        PhiloxRng r = *this;
        r.c2 = stream;
        // decorrelate repeated forks within one stage by the parent's position
        r.c1 = c1 ^ ((block * 4 - have) << 24);
        r.block = 0;
        r.have = 0;
        return r;
    }
};

}  // namespace b2pt
