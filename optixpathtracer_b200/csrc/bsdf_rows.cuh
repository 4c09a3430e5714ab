// bsdf_rows.cuh — one BSDF call per row: the per-thread bodies of the
// b2pt_bsdf_f / b2pt_bsdf_sample / b2pt_bsdf_pdf parity hooks (include/b2pt.h).
// B2PT_HD so that the CPU test tier (tests/hostsim) can run the identical code.
#pragma once

#include "../../include/b2pt.h"
#include "bsdf.cuh"
#include "fast_layered.cuh"

namespace b2pt {

// the generator of a row: the LCG policies start from the row's 32-bit state and report it back; the production
// generator is keyed the way a render keys it, with (seed, pixel, sample) = (seed[i], 0, row), and reports 0
template <class Rng>
struct RowRng {
    static B2PT_HD Rng open(uint32_t seed, int64_t) { return Rng(seed); }
    static B2PT_HD uint32_t close(const Rng& r) { return r.s; }
};
template <>
struct RowRng<PhiloxRng> {
    static B2PT_HD PhiloxRng open(uint32_t seed, int64_t i) { return PhiloxRng((uint64_t)seed, 0u, (uint32_t)i, 0u); }
    static B2PT_HD uint32_t close(const PhiloxRng&) { return 0u; }
};

// bsdf = B2PT_BSDF_CONDUCTOR .. B2PT_BSDF_LAYERED (1..4)
template <class Rng>
B2PT_HD void bsdf_row_f(int bsdf, int64_t i, const float* albedo, const float* roughness, const float* wo,
                        const float* wi, const uint32_t* seed, float* f_out, uint32_t* seed_out) {
    const ShadePoint sp = {v3(albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2]), roughness[i]};
    const float3 o = v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    const float3 in = v3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
    Rng rng = RowRng<Rng>::open(seed ? seed[i] : 0u, i);
    float3 f = v3(0.0f);
    switch (bsdf) {
        case B2PT_BSDF_CONDUCTOR: f = bsdf_f<kBsdfConductor>(rng, sp, o, in); break;
        case B2PT_BSDF_DIFFUSE: f = bsdf_f<kBsdfDiffuse>(rng, sp, o, in); break;
        case B2PT_BSDF_DIELECTRIC: f = bsdf_f<kBsdfDielectric>(rng, sp, o, in); break;
        case B2PT_BSDF_LAYERED: f = bsdf_f<kBsdfLayered>(rng, sp, o, in); break;
        default: break;
    }
    f_out[3 * i] = f.x;
    f_out[3 * i + 1] = f.y;
    f_out[3 * i + 2] = f.z;
    if (seed_out) seed_out[i] = RowRng<Rng>::close(rng);
}

template <class Rng>
B2PT_HD void bsdf_row_sample(int bsdf, int64_t i, const float* albedo, const float* roughness, const float* wo,
                             const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                             uint8_t* flags_out, uint32_t* seed_out) {
    const ShadePoint sp = {v3(albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2]), roughness[i]};
    const float3 o = v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    Rng rng = RowRng<Rng>::open(seed[i], i);
    BsdfSample s;
    s.f = s.wi = v3(0.0f);
    s.pdf = 0.0f;
    s.flags = 0;
    bool good = false;
    switch (bsdf) {
        case B2PT_BSDF_CONDUCTOR: good = bsdf_sample<kBsdfConductor>(rng, sp, o, s); break;
        case B2PT_BSDF_DIFFUSE: good = bsdf_sample<kBsdfDiffuse>(rng, sp, o, s); break;
        case B2PT_BSDF_DIELECTRIC: good = bsdf_sample<kBsdfDielectric>(rng, sp, o, s); break;
        case B2PT_BSDF_LAYERED: good = bsdf_sample<kBsdfLayered>(rng, sp, o, s); break;
        default: break;
    }
    ok[i] = good ? 1 : 0;
    f_out[3 * i] = good ? s.f.x : 0.0f;
    f_out[3 * i + 1] = good ? s.f.y : 0.0f;
    f_out[3 * i + 2] = good ? s.f.z : 0.0f;
    wi_out[3 * i] = good ? s.wi.x : 0.0f;
    wi_out[3 * i + 1] = good ? s.wi.y : 0.0f;
    wi_out[3 * i + 2] = good ? s.wi.z : 0.0f;
    pdf_out[i] = good ? s.pdf : 0.0f;
    flags_out[i] = good ? (uint8_t)s.flags : 0;
    if (seed_out) seed_out[i] = RowRng<Rng>::close(rng);
}

// ---- rows of the production (Philox) walk engine, fast_layered.cuh: one whole shade per row --------------------
// seed[i] keys the streams the way a render does with (pixel, sample) = (seed, row).
B2PT_HD fastlayer::WalkKey fast_row_key(uint32_t seed, int64_t i, uint32_t walk) {
    fastlayer::WalkKey k;
    k.k0 = seed;
    k.k1 = (uint32_t)i * 0x9E3779B9u;
    k.c1 = fastlayer::kStageWalk << 16;
    k.c2 = walk;
    return k;
}
// f = entrance reflection + (sum of the five walks) / 5; exactly 0 across the surface
B2PT_HD float3 fast_layered_f(float3 albedo, float roughness, float3 wo, float3 wi, uint32_t seed, int64_t row) {
    if (!(wo.z * wi.z > 0.0f)) return v3(0.0f);
    const float alpha = roughness * roughness;
    const float lam_wi = fastlayer::walk_lambda_wi(wi, alpha);
    float3 sum = v3(0.0f);
    for (int w = 0; w < fastlayer::kWalks; w++) {
        fastlayer::Walk k;
        fastlayer::walk_begin_f(k, wo, wi, albedo, alpha, lam_wi);
        const fastlayer::WalkKey key = fast_row_key(seed, row, (uint32_t)w);
        while (fastlayer::walk_step(k, key)) {}
        sum += k.f;
    }
    fastlayer::Walk k0;
    fastlayer::walk_begin_f(k0, wo, wi, albedo, alpha, lam_wi);
    return v3(fastlayer::entrance_reflection(k0.wo, k0.wi, k0.alpha)) + sum * (1.0f / (float)fastlayer::kWalks);
}
B2PT_HD void fast_layered_row_f(int64_t i, const float* albedo, const float* roughness, const float* wo, const float* wi,
                                const uint32_t* seed, float* f_out) {
    const float3 f = fast_layered_f(v3(albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2]), roughness[i],
                                    v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), v3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]),
                                    seed ? seed[i] : 0u, i);
    f_out[3 * i] = f.x;
    f_out[3 * i + 1] = f.y;
    f_out[3 * i + 2] = f.z;
}
// the sample walk: wi and the weight f |cos| / pdf, reported as f = weight / |wi.z|, pdf = 1
B2PT_HD void fast_layered_row_sample(int64_t i, const float* albedo, const float* roughness, const float* wo,
                                     const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                                     uint8_t* flags_out) {
    fastlayer::Walk k;
    const float3 o = v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    fastlayer::walk_begin_s(k, o, v3(albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2]), roughness[i] * roughness[i]);
    const fastlayer::WalkKey key = fast_row_key(seed[i], i, (uint32_t)fastlayer::kWalks);
    while (fastlayer::walk_step(k, key)) {}
    const bool good = k.ok;
    const float inv = good ? 1.0f / fabsf(k.w.z) : 0.0f;
    ok[i] = good ? 1 : 0;
    f_out[3 * i] = good ? k.beta.x * inv : 0.0f;
    f_out[3 * i + 1] = good ? k.beta.y * inv : 0.0f;
    f_out[3 * i + 2] = good ? k.beta.z * inv : 0.0f;
    wi_out[3 * i] = good ? k.w.x : 0.0f;
    wi_out[3 * i + 1] = good ? k.w.y : 0.0f;
    wi_out[3 * i + 2] = good ? k.w.z : 0.0f;
    pdf_out[i] = good ? 1.0f : 0.0f;
    flags_out[i] = good ? (uint8_t)(o.z * k.w.z > 0.0f ? kLobeReflection : kLobeTransmission) : 0;
}

B2PT_HD void bsdf_row_pdf(int bsdf, int64_t i, const float* roughness, const float* wo, const float* wi,
                          bool reflection, bool transmission, float* pdf_out) {
    const ShadePoint sp = {v3(1.0f), roughness[i]};
    const float3 o = v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    const float3 in = v3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
    float p = 0.0f;
    if (bsdf == B2PT_BSDF_DIELECTRIC) p = dielectric_pdf(sp, o, in, reflection, transmission);
    else if (bsdf == B2PT_BSDF_DIFFUSE) p = diffuse_pdf(o, in, reflection);
    pdf_out[i] = p;
}

}  // namespace b2pt
