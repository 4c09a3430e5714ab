// bsdf.cuh — the four BSDFs of the path-tracing core, per-thread logic (K7a-d).
//
// Re-statement for the wavefront shade kernels of the reference's pbrt-v4
// derived BxDFs (Renderer/OptiX/PBRT/*.h).  Everything works in the local
// shading frame (z = shading normal).  Functions are templated on the RNG
// policy (rng.cuh) so that the same code serves the production Philox stream
// and the two LCG flavours used for parity against the reference.
//
// Arithmetic follows the reference expression by expression (association order
// included) because the parity bar is 1e-5 relative against the reference's
// own functions compiled for the host (oracle/ref_harness.cpp); where the
// reference calls the anisotropic microfacet forms with alpha.x == alpha.y
// (Microfacet.h:22-24,66-68,81-84) the same anisotropic expressions are kept.
//
//   conductor   Conductor.h:97 (f), :122 (Sample_f), :42-95 (complex Fresnel)
//   diffuse     LambertDiffuse.h:86 (f), :110 (Sample_f), :134 (PDF)
//   dielectric  Dielectric.h:96 (f), :146 (Sample_f), :290 (PDF); eta = 1.5
//   layered     GlossyDiffuse.h:141 (f), :372 (Sample_f): pbrt LayeredBxDF,
//               dielectric over diffuse, thickness .01, 5 walks, depth 10,
//               no medium (albedo 0 => the Henyey-Greenstein code is dead)
#pragma once

#include <float.h>

#include "rng.cuh"
#include "hostmath.cuh"

namespace b2pt {

constexpr uint32_t kLobeReflection = 1u, kLobeTransmission = 2u, kLobeSpecular = 4u, kLobeGlossy = 8u;

struct BsdfSample {
    float3 f;        // BSDFSample.color
    float pdf;
    float3 wi;       // BSDFSample.direction
    uint32_t flags;  // kLobe*
};

struct ShadePoint {   // the part of `Surface` (Surface.h:5-39) the BSDFs read
    float3 albedo;
    float roughness;
};

B2PT_HD float roughness_to_alpha(float roughness) { return sqr(roughness); }   // Surface.h:26
B2PT_HD bool effectively_smooth(float alpha) { return alpha < 1e-3f; }          // Surface.h:22
B2PT_HD bool same_hemisphere(float3 a, float3 b) { return a.z * b.z > 0.0f; }
B2PT_HD bool nonzero(float3 a) { return !(a.x == 0.0f && a.y == 0.0f && a.z == 0.0f); }

// ---- spherical helpers (SphericalGeometry.h:8-35) -------------------------
B2PT_HD float cos2_theta(float3 w) { return sqr(w.z); }
B2PT_HD float sin2_theta(float3 w) { return fmaxf(0.0f, 1.0f - cos2_theta(w)); }
B2PT_HD float sin_theta(float3 w) { return sqrtf(sin2_theta(w)); }
B2PT_HD float tan2_theta(float3 w) { return sin2_theta(w) / cos2_theta(w); }
B2PT_HD float cos_phi(float3 w) {
    const float s = sin_theta(w);
    return s == 0.0f ? 1.0f : clampf(w.x / s, -1.0f, 1.0f);
}
B2PT_HD float sin_phi(float3 w) {
    const float s = sin_theta(w);
    return s == 0.0f ? 0.0f : clampf(w.y / s, -1.0f, 1.0f);
}

// ---- Trowbridge-Reitz microfacet model (Microfacet.h) ---------------------
B2PT_HD_CALL float ggx_d(float3 wm, float alpha) {                       // :9-20
    const float t2 = tan2_theta(wm);
    if (isinf(t2)) return 0.0f;
    const float c4 = sqr(cos2_theta(wm));
    if (c4 < 1e-16f) return 0.0f;
    const float e = t2 * (sqr(cos_phi(wm) / alpha) + sqr(sin_phi(wm) / alpha));
    return 1.0f / (kPi * alpha * alpha * c4 * sqr(1.0f + e));
}
B2PT_HD_CALL float ggx_lambda(float3 w, float alpha) {                   // :46-52
    const float t2 = tan2_theta(w);
    if (isinf(t2)) return 0.0f;
    const float a2 = sqr(cos_phi(w) * alpha) + sqr(sin_phi(w) * alpha);
    return (sqrtf(1.0f + a2 * t2) - 1.0f) / 2.0f;
}
B2PT_HD float ggx_g(float3 wo, float3 wi, float alpha) {            // :62-64
    return 1.0f / (1.0f + ggx_lambda(wo, alpha) + ggx_lambda(wi, alpha));
}
B2PT_HD float ggx_g1(float3 w, float alpha) { return 1.0f / (1.0f + ggx_lambda(w, alpha)); }  // :71-73
// distribution of visible normals = the sampling pdf of sample_wm (:81-88)
B2PT_HD float ggx_visible_pdf(float3 w, float3 wm, float alpha) {
    return ggx_g1(w, alpha) / fabsf(w.z) * ggx_d(wm, alpha) * absdot(w, wm);
}

// Heitz visible-normal sampling (:90-121); two draws, taken as the pair
// `glm::vec2(rnd, rnd)` of SampleUniformDiskPolar (random.h:76-84).
// The body is a real call on the device (B2PT_HD_CALL): the GGX helpers are reached from many places
// of the dielectric / layered code and, inlined everywhere, made kernels of 7 000 - 18 000 instructions
// that stalled on instruction fetch.
B2PT_HD_CALL float3 ggx_sample_wm_u(float3 w, float alpha, float ux, float uy) {
    float3 wh = normalize(v3(alpha * w.x, alpha * w.y, w.z));
    if (wh.z < 0.0f) wh = -wh;
    const float3 t1 = wh.z < 0.99999f ? normalize(cross(v3(0.0f, 0.0f, 1.0f), wh)) : v3(1.0f, 0.0f, 0.0f);
    const float3 t2 = cross(wh, t1);
    const float r = sqrtf(ux);
    const float theta = 2.0f * kPi * uy;
    float px = r * hostmath::cosf_host(theta), py = r * hostmath::sinf_host(theta);
    const float h = sqrtf(1.0f - sqr(px));
    const float lerp_t = (1.0f + wh.z) / 2.0f;
    py = (1.0f - lerp_t) * h + lerp_t * py;
    const float pz = sqrtf(fmaxf(0.0f, 1.0f - (sqr(px) + sqr(py))));
    const float3 nh = px * t1 + py * t2 + pz * wh;
    return normalize(v3(alpha * nh.x, alpha * nh.y, fmaxf(1e-6f, nh.z)));
}
template <class Rng>
B2PT_HD float3 ggx_sample_wm(Rng& rng, float3 w, float alpha) {
    float ux, uy;
    rng.next2(ux, uy);
    return ggx_sample_wm_u(w, alpha, ux, uy);
}

// ---- Fresnel ---------------------------------------------------------------
// unpolarised dielectric Fresnel (Dielectric.h:20-44)
B2PT_HD_CALL float fresnel_dielectric(float cos_i, float ior) {
    cos_i = clampf(cos_i, -1.0f, 1.0f);
    if (cos_i < 0.0f) {
        ior = 1.0f / ior;
        cos_i = -cos_i;
    }
    const float s2i = 1.0f - sqr(cos_i);
    const float s2t = s2i / sqr(ior);
    if (s2t >= 1.0f) return 1.0f;
    const float cos_t = sqrtf(1.0f - s2t);
    const float r_parl = (ior * cos_i - cos_t) / (ior * cos_i + cos_t);
    const float r_perp = (cos_i - ior * cos_t) / (cos_i + ior * cos_t);
    return (sqr(r_parl) + sqr(r_perp)) / 2.0f;
}

// complex arithmetic with the operation order of Complex.h:5-63
struct Cplx {
    float re, im;
};
B2PT_HD Cplx cmul(Cplx a, Cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
B2PT_HD Cplx cdiv(Cplx a, Cplx b) {
    const float scale = 1.0f / (b.re * b.re + b.im * b.im);
    return {scale * (a.re * b.re + a.im * b.im), scale * (a.im * b.re - a.re * b.im)};
}
B2PT_HD Cplx cadd(Cplx a, Cplx b) { return {a.re + b.re, a.im + b.im}; }
B2PT_HD Cplx csub(Cplx a, Cplx b) { return {a.re - b.re, a.im - b.im}; }
B2PT_HD float cnorm(Cplx a) { return a.re * a.re + a.im * a.im; }
B2PT_HD Cplx csqrt(Cplx z) {
    const float n = sqrtf(cnorm(z));
    const float t1 = sqrtf(0.5f * (n + fabsf(z.re)));
    const float t2 = 0.5f * z.im / t1;
    if (n == 0.0f) return {0.0f, 0.0f};
    if (z.re >= 0.0f) return {t1, t2};
    return {fabsf(t2), copysignf(t1, z.im)};
}
// Fresnel reflectance of a conductor with complex IOR eta + i k (Conductor.h:42-52)
B2PT_HD_CALL float fresnel_complex_1(float cos_i, Cplx eta) {
    cos_i = clampf(cos_i, 0.0f, 1.0f);
    const Cplx ci = {cos_i, 0.0f};
    const float s2i = 1.0f - sqr(cos_i);
    const Cplx s2t = cdiv({s2i, 0.0f}, cmul(eta, eta));
    const Cplx ct = csqrt(csub({1.0f, 0.0f}, s2t));
    const Cplx eci = cmul(eta, ci), ect = cmul(eta, ct);
    const Cplx r_parl = cdiv(csub(eci, ct), cadd(eci, ct));
    const Cplx r_perp = cdiv(csub(ci, ect), cadd(ci, ect));
    return (cnorm(r_parl) + cnorm(r_perp)) / 2.0f;
}
// per channel, IOR derived from the reflectance colour (Conductor.h:54-93):
// eta = 1, k = 2 sqrt(r) / sqrt(1 - r), r clamped to [0, .9999]
B2PT_HD_CALL float3 fresnel_conductor(float cos_i, float3 reflectance) {
    const float3 r = v3(clampf(reflectance.x, 0.0f, 0.9999f), clampf(reflectance.y, 0.0f, 0.9999f),
                        clampf(reflectance.z, 0.0f, 0.9999f));
    const float kx = 2.0f * sqrtf(r.x) / sqrtf(fmaxf(1.0f - r.x, 0.0f));
    const float ky = 2.0f * sqrtf(r.y) / sqrtf(fmaxf(1.0f - r.y, 0.0f));
    const float kz = 2.0f * sqrtf(r.z) / sqrtf(fmaxf(1.0f - r.z, 0.0f));
    return v3(fresnel_complex_1(cos_i, {1.0f, kx}), fresnel_complex_1(cos_i, {1.0f, ky}),
              fresnel_complex_1(cos_i, {1.0f, kz}));
}

// ===========================================================================
// K7a — conductor (Conductor.h:97-120, :122-196)
// ===========================================================================
B2PT_HD float3 conductor_f(const ShadePoint& sp, float3 wo, float3 wi) {
    const float alpha = roughness_to_alpha(sp.roughness);
    if (!same_hemisphere(wo, wi)) return v3(0.0f);
    if (effectively_smooth(alpha)) return v3(0.0f);
    const float cos_o = fabsf(wo.z), cos_i = fabsf(wi.z);
    if (cos_i == 0.0f || cos_o == 0.0f) return v3(0.0f);
    float3 wm = wi + wo;
    if (length_sqr(wm) == 0.0f) return v3(0.0f);
    wm = normalize(wm);
    const float3 F = fresnel_conductor(absdot(wo, wm), sp.albedo);
    return ggx_d(wm, alpha) * F * ggx_g(wo, wi, alpha) / (4.0f * cos_i * cos_o);
}

template <class Rng>
B2PT_HD bool conductor_sample(Rng& rng, const ShadePoint& sp, float3 wo, BsdfSample& s) {
    const float alpha = roughness_to_alpha(sp.roughness);
    if (effectively_smooth(alpha)) {  // perfect mirror, no draws
        const float3 wi = v3(-wo.x, -wo.y, wo.z);
        s.f = fresnel_conductor(fabsf(wi.z), sp.albedo) / fabsf(wi.z);
        s.wi = wi;
        s.pdf = 1.0f;
        s.flags = kLobeReflection | kLobeSpecular;
        return true;
    }
    if (wo.z == 0.0f) return false;
    const float3 wm = ggx_sample_wm(rng, wo, alpha);
    const float3 wi = -wo + 2.0f * dot(wo, wm) * wm;
    if (!same_hemisphere(wo, wi)) return false;
    const float pdf = ggx_visible_pdf(wo, wm, alpha) / (4.0f * absdot(wo, wm));
    const float cos_o = fabsf(wo.z), cos_i = fabsf(wi.z);
    if (cos_i == 0.0f || cos_o == 0.0f) return false;
    const float3 F = fresnel_conductor(absdot(wo, wm), sp.albedo);
    s.f = ggx_d(wm, alpha) * F * ggx_g(wo, wi, alpha) / (4.0f * cos_i * cos_o);
    s.wi = wi;
    s.pdf = pdf;
    s.flags = kLobeReflection | kLobeGlossy;
    return true;
}

// ===========================================================================
// K7b — Lambert diffuse (LambertDiffuse.h:86-139)
// ===========================================================================
B2PT_HD float3 diffuse_f(const ShadePoint& sp, float3 wo, float3 wi) {
    if (!same_hemisphere(wo, wi)) return v3(0.0f);
    return sp.albedo * kInvPi;
}

// concentric disk map + cosine lift (:35-61); the two draws are one
// `glm::vec2(rnd, rnd)` pair
template <class Rng>
B2PT_HD float3 sample_cosine_hemisphere(Rng& rng) {
    const float pi_over_4 = 0.78539816339744830961f, pi_over_2 = 1.57079632679489661923f;
    float ux, uy;
    rng.next2(ux, uy);
    const float ox = 2.0f * ux - 1.0f, oy = 2.0f * uy - 1.0f;
    float dx = 0.0f, dy = 0.0f;
    if (!(ox == 0.0f && oy == 0.0f)) {
        float theta, r;
        if (fabsf(ox) > fabsf(oy)) {
            r = ox;
            theta = pi_over_4 * (oy / ox);
        } else {
            r = oy;
            theta = pi_over_2 - pi_over_4 * (ox / oy);
        }
        dx = r * hostmath::cosf_host(theta);
        dy = r * hostmath::sinf_host(theta);
    }
    const float z = sqrtf(fmaxf(0.0f, 1.0f - sqr(dx) - sqr(dy)));
    return v3(dx, dy, z);
}

// `reflection` = the lobe switch of Layer_Sample_f; without it nothing is drawn.
// The sample is never flipped to wo's side.
template <class Rng>
B2PT_HD bool diffuse_sample(Rng& rng, const ShadePoint& sp, float3 /*wo*/, bool reflection, BsdfSample& s) {
    if (!reflection) return false;
    float3 wi = sample_cosine_hemisphere(rng);
    if (wi.z < 0.0f) wi.z *= -1.0f;
    wi = normalize(wi);
    s.wi = wi;
    s.pdf = fabsf(wi.z) * kInvPi;
    s.f = sp.albedo * kInvPi;
    s.flags = kLobeReflection;
    return true;
}

B2PT_HD float diffuse_pdf(float3 wo, float3 wi, bool reflection) {
    if (!reflection || !same_hemisphere(wi, wo)) return 0.0f;
    return fabsf(wi.z) * kInvPi;
}

// ===========================================================================
// K7c — dielectric, eta = 1.5 (Dielectric.h)
// ===========================================================================
constexpr float kDielectricEta = 1.5f;

// Snell refraction about n (:68-92); false on total internal reflection.
B2PT_HD bool refract(float3 wi, float3 n, float eta, float& etap, float3& wt) {
    float cos_i = dot(n, wi);
    if (cos_i < 0.0f) {
        eta = 1.0f / eta;
        cos_i = -cos_i;
        n = -n;
    }
    const float s2i = fmaxf(0.0f, 1.0f - sqr(cos_i));
    const float s2t = s2i / sqr(eta);
    if (s2t >= 1.0f) return false;
    const float cos_t = sqrtf(1.0f - s2t);
    wt = -wi / eta + (cos_i / eta - cos_t) * n;
    etap = eta;
    return true;
}

// generalised half vector, oriented to +z: glm::faceforward(-n, (0,0,1), n)
B2PT_HD_CALL float3 half_vector_up(float3 wm) {
    const float3 n = normalize(wm);
    return n.z < 0.0f ? -n : n;
}

// radiance = TransportMode::Radiance (divide transmission by etap^2)
B2PT_HD float3 dielectric_f(const ShadePoint& sp, float3 wo, float3 wi, bool radiance) {     // :96-139
    const float eta = kDielectricEta;
    const float alpha = roughness_to_alpha(sp.roughness);
    if (effectively_smooth(alpha)) return v3(0.0f);
    const float cos_o = wo.z, cos_i = wi.z;
    const bool reflect = cos_i * cos_o > 0.0f;
    float etap = 1.0f;
    if (!reflect) etap = cos_o > 0.0f ? eta : (1.0f / eta);
    float3 wm = wi * etap + wo;
    if (cos_i == 0.0f || cos_o == 0.0f || sqr(length(wm)) == 0.0f) return v3(0.0f);
    wm = half_vector_up(wm);
    if (dot(wm, wi) * cos_i < 0.0f || dot(wm, wo) * cos_o < 0.0f) return v3(0.0f);
    const float F = fresnel_dielectric(dot(wo, wm), eta);
    if (reflect) {
        return v3(ggx_d(wm, alpha) * ggx_g(wo, wi, alpha) * F / fabsf(4.0f * cos_i * cos_o));
    }
    const float denom = sqr(dot(wi, wm) + dot(wo, wm) / etap) * cos_i * cos_o;
    float ft = ggx_d(wm, alpha) * (1.0f - F) * ggx_g(wo, wi, alpha) * fabsf(dot(wi, wm) * dot(wo, wm) / denom);
    if (radiance) ft /= sqr(etap);
    return v3(ft);
}

// (:146-288) uc is drawn first in every case, then the two microfacet draws.
template <class Rng>
B2PT_HD bool dielectric_sample(Rng& rng, const ShadePoint& sp, float3 wo, bool radiance, bool reflection,
                               bool transmission, BsdfSample& s) {
    const float eta = kDielectricEta;
    const float alpha = roughness_to_alpha(sp.roughness);
    const float uc = rng.next();
    if (effectively_smooth(alpha)) {
        const float R = fresnel_dielectric(wo.z, eta), T = 1.0f - R;
        float pr = R, pt = T;
        if (!reflection) pr = 0.0f;
        if (!transmission) pt = 0.0f;
        if (pr == 0.0f && pt == 0.0f) return false;
        if (uc < pr / (pr + pt)) {
            const float3 wi = v3(-wo.x, -wo.y, wo.z);
            s.f = v3(R / fabsf(wi.z));
            s.wi = wi;
            s.pdf = pr / (pr + pt);
            s.flags = kLobeReflection | kLobeSpecular;
            return true;
        }
        float3 wi;
        float etap;
        if (!refract(wo, v3(0.0f, 0.0f, 1.0f), eta, etap, wi)) return false;
        float ft = T / fabsf(wi.z);
        if (radiance) ft /= sqr(etap);
        s.f = v3(ft);
        s.wi = wi;
        s.pdf = pt / (pr + pt);
        s.flags = kLobeTransmission | kLobeSpecular;
        return true;
    }
    const float3 wm = ggx_sample_wm(rng, wo, alpha);
    const float R = fresnel_dielectric(dot(wo, wm), eta), T = 1.0f - R;
    float pr = R, pt = T;
    if (!reflection) pr = 0.0f;
    if (!transmission) pt = 0.0f;
    if (pr == 0.0f && pt == 0.0f) return false;
    if (uc < pr / (pr + pt)) {
        // glm::reflect(-wo, wm) = I - N * dot(N, I) * 2 with I = -wo  
        const float3 I = -wo;
        const float3 wi = I - wm * dot(wm, I) * 2.0f;
        if (!same_hemisphere(wo, wi)) return false;
        s.pdf = ggx_visible_pdf(wo, wm, alpha) / (4.0f * absdot(wo, wm)) * pr / (pr + pt);
        s.f = v3(ggx_d(wm, alpha) * ggx_g(wo, wi, alpha) * R / (4.0f * wi.z * wo.z));
        s.wi = wi;
        s.flags = kLobeReflection | kLobeGlossy;
        return true;
    }
    float etap = 0.0f;
    float3 wi = v3(0.0f);
    const bool tir = !refract(wo, wm, eta, etap, wi);
    if (tir || same_hemisphere(wo, wi) || wi.z == 0.0f) return false;
    const float denom = sqr(dot(wi, wm) + dot(wo, wm) / etap);
    const float dwm_dwi = absdot(wi, wm) / denom;
    s.pdf = ggx_visible_pdf(wo, wm, alpha) * dwm_dwi * pt / (pr + pt);
    float ft = T * ggx_d(wm, alpha) * ggx_g(wo, wi, alpha) *
               fabsf(dot(wi, wm) * dot(wo, wm) / (wi.z * wo.z * denom));
    if (radiance) ft /= sqr(etap);
    s.f = v3(ft);
    s.wi = wi;
    s.flags = kLobeTransmission | kLobeGlossy;
    return true;
}

B2PT_HD float dielectric_pdf(const ShadePoint& sp, float3 wo, float3 wi, bool reflection, bool transmission) {  // :290-341
    const float eta = kDielectricEta;
    const float alpha = roughness_to_alpha(sp.roughness);
    if (effectively_smooth(alpha)) return 0.0f;
    const float cos_o = wo.z, cos_i = wi.z;
    const bool reflect = cos_i * cos_o > 0.0f;
    float etap = 1.0f;
    if (!reflect) etap = cos_o > 0.0f ? eta : (1.0f / eta);
    float3 wm = wi * etap + wo;
    if (cos_i == 0.0f || cos_o == 0.0f || length_sqr(wm) == 0.0f) return 0.0f;
    wm = half_vector_up(wm);
    if (dot(wm, wi) * cos_i < 0.0f || dot(wm, wo) * cos_o < 0.0f) return 0.0f;
    const float R = fresnel_dielectric(dot(wo, wm), eta), T = 1.0f - R;
    float pr = R, pt = T;
    if (!reflection) pr = 0.0f;
    if (!transmission) pt = 0.0f;
    if (pr == 0.0f && pt == 0.0f) return 0.0f;
    if (reflect) return ggx_visible_pdf(wo, wm, alpha) / (4.0f * absdot(wo, wm)) * pr / (pr + pt);
    const float denom = sqr(dot(wi, wm) + dot(wo, wm) / etap);
    const float dwm_dwi = absdot(wi, wm) / denom;
    return ggx_visible_pdf(wo, wm, alpha) * dwm_dwi * pt / (pr + pt);
}

// ===========================================================================
// K7d — layered: dielectric coating over a diffuse base (GlossyDiffuse.h)
//
// pbrt-v4's LayeredBxDF<Dielectric, Diffuse, twoSided = true> evaluated by
// stochastic random walks between the two interfaces.  The reference fixes
// mediaAlbedo = 0 and g = 0 (:148-151, :378-381), so a walk never scatters
// between the interfaces: it only bounces top <-> bottom, attenuated by
// exp(-thickness/|w.z|), and the phase-function branches (:269-312, :443-472)
// are unreachable; they are not built here.
//
// Which generator feeds what: every interface sample draws from
// the PATH stream `rng`; only the Russian-roulette test draws from a side
// stream forked from (wo, wi, current path state) (:215-222, :417-422).
// ===========================================================================
constexpr float kLayerThickness = 0.01f;
constexpr int kLayerWalks = 5;       // nSamples
constexpr int kLayerMaxDepth = 10;

B2PT_HD_CALL float layer_transmittance(float dz, float3 w) {                 // :97-106
    if (fabsf(dz) <= FLT_MIN) return 1.0f;
    return hostmath::expf_host(-fabsf(dz / w.z));
}
// float3 / float as a real call on the device: the walks divide by a pdf in a dozen places and each
// inline IEEE division is ~12 instructions x 3 components
B2PT_HD_CALL float3 div3(float3 a, float s) { return a / s; }
B2PT_HD float power_heuristic_1(float fpdf, float gpdf) {               // :91-95, nf = ng = 1
    const float f = 1.0f * fpdf, g = 1.0f * gpdf;
    return sqr(f) / (sqr(f) + sqr(g));
}

// the two interfaces behind one switch (Layer_f / Layer_Sample_f / Layer_PDF, :110-135).
// Kept as real calls on the device: the walks below reach them from a dozen places, and inlined the
// layered kernel was 18 000 instructions (290 KB) whose warps stalled on instruction fetch 70 % of
// the time (profiles/r01_k_shade_layered_*.csv).
B2PT_HD_CALL float3 interface_f(bool top, const ShadePoint& sp, float3 wo, float3 wi, bool radiance) {
    return top ? dielectric_f(sp, wo, wi, radiance) : diffuse_f(sp, wo, wi);
}
template <class Rng>
B2PT_HD_CALL bool interface_sample(bool top, Rng& rng, const ShadePoint& sp, float3 wo, bool radiance,
                              bool reflection, bool transmission, BsdfSample& s) {
    return top ? dielectric_sample(rng, sp, wo, radiance, reflection, transmission, s)
               : diffuse_sample(rng, sp, wo, reflection, s);
}
B2PT_HD_CALL float interface_pdf(bool top, const ShadePoint& sp, float3 wo, float3 wi, bool reflection,
                            bool transmission) {
    return top ? dielectric_pdf(sp, wo, wi, reflection, transmission) : diffuse_pdf(wo, wi, reflection);
}
B2PT_HD bool usable(bool ok, const BsdfSample& s) {   // the reference's `!test || color == 0 || pdf == 0 || dir.z == 0`
    return ok && nonzero(s.f) && s.pdf != 0.0f && s.wi.z != 0.0f;
}

// GlossyDiffuse::f (:141-367)
template <class Rng>
B2PT_HD float3 layered_f(Rng& rng, const ShadePoint& sp, float3 wo, float3 wi) {
    const float thickness = kLayerThickness;
    if (wo.z < 0.0f) {  // two-sided
        wo = -wo;
        wi = -wi;
    }
    const bool top_specular = effectively_smooth(roughness_to_alpha(sp.roughness));
    const bool same = same_hemisphere(wo, wi);
    // entered through the top; the walk must leave through the top when wi is on
    // wo's side and through the (opaque, diffuse) bottom otherwise
    const bool exit_top = same;
    const bool nonexit_top = !same;
    const bool exit_specular = exit_top ? top_specular : false;
    const bool nonexit_specular = nonexit_top ? top_specular : false;
    const float exit_z = same ? thickness : 0.0f;

    float3 f = v3(0.0f);
    if (same) f = v3((float)kLayerWalks) * interface_f(true, sp, wo, wi, true);

    Rng rr = rng.fork4(wo, wi);

    for (int s = 0; s < kLayerWalks; ++s) {
        BsdfSample wos;
        const bool wos_ok = interface_sample(true, rng, sp, wo, true, false, true, wos);
        if (!usable(wos_ok, wos)) continue;
        BsdfSample wis;  // virtual light direction through the exit interface, importance transport
        const bool wis_ok = interface_sample(exit_top, rng, sp, wi, false, false, true, wis);
        if (!usable(wis_ok, wis)) continue;

        float3 beta = div3(wos.f * fabsf(wos.wi.z), wos.pdf);
        float z = thickness;
        float3 w = wos.wi;
        for (int depth = 0; depth < kLayerMaxDepth; ++depth) {
            if (depth > 3 && max3(beta) < 0.25f) {
                const float q = fmaxf(0.0f, 1.0f - max3(beta));
                if (rr.next() < q) break;
                beta = div3(beta, 1.0f - q);
            }
            z = (z == thickness) ? 0.0f : thickness;
            beta *= layer_transmittance(thickness, w);
            if (z == exit_z) {
                BsdfSample bs;
                const bool ok = interface_sample(exit_top, rng, sp, -w, true, true, false, bs);
                if (!usable(ok, bs)) break;
                beta *= div3(bs.f * fabsf(bs.wi.z), bs.pdf);
                w = bs.wi;
            } else {
                if (!nonexit_specular) {  // NEE along the presampled wis
                    float wt = 1.0f;
                    if (!exit_specular)
                        wt = power_heuristic_1(wis.pdf, interface_pdf(nonexit_top, sp, -w, -wis.wi, true, true));
                    f += div3(beta * interface_f(nonexit_top, sp, -w, -wis.wi, true) * fabsf(wis.wi.z) * wt *
                                  layer_transmittance(thickness, wis.wi) * wis.f, wis.pdf);
                }
                BsdfSample bs;
                const bool ok = interface_sample(nonexit_top, rng, sp, -w, true, true, false, bs);
                if (!usable(ok, bs)) break;
                beta *= div3(bs.f * fabsf(bs.wi.z), bs.pdf);
                w = bs.wi;
                if (!exit_specular) {  // NEE along the sampled direction
                    const float3 f_exit = interface_f(exit_top, sp, -w, wi, true);
                    if (nonzero(f_exit)) {
                        float wt = 1.0f;
                        if (!nonexit_specular) {
                            const float exit_pdf = interface_pdf(exit_top, sp, -w, wi, false, true);
                            wt = power_heuristic_1(bs.pdf, exit_pdf);
                        }
                        f += beta * layer_transmittance(thickness, bs.wi) * f_exit * wt;
                    }
                }
            }
        }
    }
    return f / v3((float)kLayerWalks);
}

// GlossyDiffuse::Sample_f (:372-514); the returned pdf is pbrt's
// "pdfIsProportional" estimate
template <class Rng>
B2PT_HD bool layered_sample(Rng& rng, const ShadePoint& sp, float3 wo, BsdfSample& out) {
    const float thickness = kLayerThickness;
    bool flip = false;
    if (wo.z < 0.0f) {
        wo = -wo;
        flip = true;
    }
    BsdfSample bs;
    const bool ok0 = interface_sample(true, rng, sp, wo, true, true, true, bs);
    if (!usable(ok0, bs)) return false;
    if (bs.flags & kLobeReflection) {
        if (flip) bs.wi = -bs.wi;
        out = bs;
        return true;
    }
    float3 w = bs.wi;
    bool specular_path = (bs.flags & kLobeSpecular) != 0;

    Rng rr = rng.fork2(wo);

    float3 f = bs.f * fabsf(bs.wi.z);
    float pdf = bs.pdf;
    float z = thickness;
    for (int depth = 0; depth < kLayerMaxDepth; ++depth) {
        const float rr_beta = max3(f) / pdf;
        if (depth > 3 && rr_beta < 0.25f) {
            const float q = fmaxf(0.0f, 1.0f - rr_beta);
            if (rr.next() < q) return false;
            pdf *= 1.0f - q;
        }
        if (w.z == 0.0f) return false;
        z = (z == thickness) ? 0.0f : thickness;
        f *= layer_transmittance(thickness, w);
        const bool at_top = !(z == 0.0f);
        BsdfSample is;
        const bool ok = interface_sample(at_top, rng, sp, -w, true, true, true, is);
        if (!usable(ok, is)) return false;
        f *= is.f;
        pdf *= is.pdf;
        specular_path = specular_path && (is.flags & kLobeSpecular) != 0;
        w = is.wi;
        if (is.flags & kLobeTransmission) {  // left the layers
            if (flip) w = -w;
            out.f = f;
            out.wi = w;
            out.pdf = pdf;
            const bool refl = same_hemisphere(wo, w);
            out.flags = (refl ? kLobeReflection : kLobeTransmission) | (specular_path ? kLobeSpecular : kLobeGlossy);
            return true;
        }
        f *= fabsf(is.wi.z);
    }
    return false;
}

// ---------------------------------------------------------------------------
// dispatch by material class (B2PT_BSDF_* - 1)
// ---------------------------------------------------------------------------
enum BsdfKind : int { kBsdfConductor = 0, kBsdfDiffuse = 1, kBsdfDielectric = 2, kBsdfLayered = 3 };

template <int KIND, class Rng>
B2PT_HD float3 bsdf_f(Rng& rng, const ShadePoint& sp, float3 wo, float3 wi) {
    if (KIND == kBsdfConductor) return conductor_f(sp, wo, wi);
    if (KIND == kBsdfDiffuse) return diffuse_f(sp, wo, wi);
    if (KIND == kBsdfDielectric) return dielectric_f(sp, wo, wi, true);
    return layered_f(rng, sp, wo, wi);
}
template <int KIND, class Rng>
B2PT_HD bool bsdf_sample(Rng& rng, const ShadePoint& sp, float3 wo, BsdfSample& s) {
    if (KIND == kBsdfConductor) return conductor_sample(rng, sp, wo, s);
    if (KIND == kBsdfDiffuse) return diffuse_sample(rng, sp, wo, true, s);
    if (KIND == kBsdfDielectric) return dielectric_sample(rng, sp, wo, true, true, true, s);
    return layered_sample(rng, sp, wo, s);
}

}  // namespace b2pt
