// bsdf_rows.cuh — one BSDF call per row: the per-thread bodies of the
// b2pt_bsdf_f / b2pt_bsdf_sample / b2pt_bsdf_pdf parity hooks (include/b2pt.h).
// B2PT_HD so that the CPU test tier (tests/hostsim) can run the identical code.
#pragma once

#include "../../include/b2pt.h"
#include "bsdf.cuh"

namespace b2pt {

// bsdf = B2PT_BSDF_CONDUCTOR .. B2PT_BSDF_LAYERED (1..4)
template <class Rng>
B2PT_HD void bsdf_row_f(int bsdf, int64_t i, const float* albedo, const float* roughness, const float* wo,
                        const float* wi, const uint32_t* seed, float* f_out, uint32_t* seed_out) {
    const ShadePoint sp = {v3(albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2]), roughness[i]};
    const float3 o = v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    const float3 in = v3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
    Rng rng(seed ? seed[i] : 0u);
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
    if (seed_out) seed_out[i] = rng.s;
}

template <class Rng>
B2PT_HD void bsdf_row_sample(int bsdf, int64_t i, const float* albedo, const float* roughness, const float* wo,
                             const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                             uint8_t* flags_out, uint32_t* seed_out) {
    const ShadePoint sp = {v3(albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2]), roughness[i]};
    const float3 o = v3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    Rng rng(seed[i]);
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
    if (seed_out) seed_out[i] = rng.s;
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
