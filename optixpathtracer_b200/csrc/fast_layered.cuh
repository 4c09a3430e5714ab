// fast_layered.cuh — the layered BSDF (dielectric coating over a diffuse base) as a
// walk engine: per-lane logic of the production (Philox) shade path, K7d.
//
// What it computes is the estimator of the reference's GlossyDiffuse
// (PBRT/GlossyDiffuse.h:141-367 `f`, :372-514 `Sample_f` — pbrt-v4's
// LayeredBxDF<Dielectric, Diffuse, twoSided>, thickness .01, 5 walks, depth 10,
// no medium), the same one bsdf.cuh restates operation by operation for the
// bit-exact LCG path.  Here the SAME estimator is re-organised for the GPU:
//
//  * A shade is cut into independent WALK tasks — five `f` walks per hit whose
//    light is visible, one `Sample_f` walk per hit — so a warp runs 32 walks,
//    not 32 hits x 6 serial walks; every task draws from its own Philox stream
//    (key = seed, pixel, sample; counter = bounce, walk, step), so the result
//    does not depend on which lane ran what.
//  * A walk is a sequence of STEPS of one shape: [event at the top interface:
//    one visible-normal GGX sample] then [event at the bottom interface: one
//    cosine sample + the two next-event terms].  Both walk kinds and the two
//    pre-samples of an `f` walk (wos from wo, wis from wi) fit that shape, so
//    the whole engine is one loop around ~1 000 instructions that stay in the
//    instruction cache (the serial kernel was 5 000 and stalled on fetch).
//  * Only sample WEIGHTS f |cos| / pdf are formed.  For the visible-normal
//    sampler the microfacet density cancels analytically:
//        reflection    f |cos_i| / pdf = G(wo, wi) / G1(wo) * (pr + pt)
//        transmission  f |cos_i| / pdf = G(wo, wi) / G1(wo) * (pr + pt) / etap^2   (radiance)
//    (Dielectric.h:226-285: pdf = D_w(wm) / (4 |wo.wm|) * pr / (pr + pt), f = D G R / (4 cos_i cos_o), and
//    the transmission pair with the half-vector Jacobian); a cosine sample of the Lambert base weighs
//    exactly `albedo`.  D is evaluated only where a pdf is needed on its own (the MIS weights).
//  * The last top event of an `f` walk (depth 9) cannot contribute any more and is not run; `f` is not
//    evaluated at all when wo and wi lie on opposite sides (every walk ends at the opaque base: exactly 0).
//
// Arithmetic: plain fp32 with FMA contraction, approximate division / square root / exp / sin / cos
// (the translation unit is built with --use_fast_math); it is NOT bit-compatible with the reference —
// the bit-exact statement of the same functions is bsdf.cuh, which the LCG modes run.  Checked
// statistically against the reference's own code (tests/test_fast_layered_*.py).
#pragma once

#include "rng.cuh"

namespace b2pt {
namespace fastlayer {

constexpr float kEta = 1.5f;             // Dielectric.h: eta of the coating
constexpr float kThickness = 0.01f;      // GlossyDiffuse.h:144
constexpr int kWalks = 5;                // nSamples :146
constexpr int kStepsF = 6;               // steps of an f walk: wos | wis + depth 0 | depths 1,2 | 3,4 | 5,6 | 7,8
constexpr int kStepsS = 6;               // steps of a sample walk: first sample + depth 0 | depths 1,2 | ... | depth 9

B2PT_HD float frsqrt(float x) {
#if defined(__CUDA_ARCH__)
    return rsqrtf(x);
#else
    return 1.0f / sqrtf(x);
#endif
}
B2PT_HD float3 fnormalize(float3 a) { return a * frsqrt(dot(a, a)); }
B2PT_HD void fsincos(float x, float& s, float& c) {
#if defined(__CUDA_ARCH__)
    __sincosf(x, &s, &c);
#else
    s = sinf(x);
    c = cosf(x);
#endif
}
B2PT_HD float fexp(float x) {
#if defined(__CUDA_ARCH__)
    return __expf(x);
#else
    return expf(x);
#endif
}

// five 24-bit uniforms from one Philox block (the reference's rnd has 24 bits too, random.h:66-69)
struct Draws {
    float uc, u1, u2, d1, d2;
};
B2PT_HD Draws draws_from_block(uint4 o) {
    const float s = 1.0f / 16777216.0f;
    Draws d;
    d.uc = (float)(o.x & 0xffffffu) * s;
    d.u1 = (float)(o.y & 0xffffffu) * s;
    d.u2 = (float)(o.z & 0xffffffu) * s;
    d.d1 = (float)(o.w & 0xffffffu) * s;
    d.d2 = (float)((o.x >> 24) | ((o.y >> 24) << 8) | ((o.z >> 24) << 16)) * s;
    return d;
}

// stream of one walk task
struct WalkKey {
    uint32_t k0, k1;   // seed ^ pixel, seed_hi ^ sample index
    uint32_t c1;       // bounce | kStageWalk << 16
    uint32_t c2;       // walk id: 0..4 f walks, 5 the sample walk
};
constexpr uint32_t kStageWalk = 2;   // stage 0 = K6 surface, 1 = serial shade kernels
// RNG = 0: the out-of-line ten-round block every other draw site calls; 10 / 7: the block inlined here with that many
// rounds (the engine's A/B, B2PT_WALK_RNG: the Philox rounds are 15 % of the engine's instructions)
template <int RNG>
B2PT_HD Draws walk_draws(const WalkKey& k, int step) {
    if constexpr (RNG == 0) return draws_from_block(philox_block((uint32_t)step, k.c1, k.c2, k.k0, k.k1));
    else return draws_from_block(philox_block_inline<RNG>((uint32_t)step, k.c1, k.c2, k.k0, k.k1));
}
// Russian-roulette uniforms are needed rarely (depth > 3 and throughput < 1/4): drawn on demand from a side block
B2PT_HD float walk_rr_uniform(const WalkKey& k, int step, int which) {
    const uint4 o = philox_block((uint32_t)step | 0x100u, k.c1, k.c2, k.k0, k.k1);
    return (float)((which ? o.y : o.x) >> 8) * (1.0f / 16777216.0f);
}

// ---- Trowbridge-Reitz, isotropic (Microfacet.h with alpha.x == alpha.y) ------------------------------
// 1 + Lambda(w)  (:46-52; Lambda = 0 where tan^2 is infinite)
B2PT_HD float lambda1p(float3 w, float a2) {
    const float c2 = w.z * w.z;
    if (c2 == 0.0f) return 1.0f;
    const float s2 = fmaxf(0.0f, 1.0f - c2);
    return 0.5f + 0.5f * sqrtf(1.0f + a2 * s2 / c2);
}
// D(wm) = a^2 / (pi (sin^2 + a^2 cos^2)^2)  (:9-20 with both alphas equal)
B2PT_HD float ggx_d(float3 wm, float a2) {
    const float c2 = wm.z * wm.z;
    if (c2 * c2 < 1e-16f) return 0.0f;
    const float t = (wm.x * wm.x + wm.y * wm.y) + a2 * c2;
    return a2 / (kPi * t * t);
}
// Heitz' visible-normal sample (:90-121), wm.z > 0
B2PT_HD float3 sample_wm(float3 w, float alpha, float u1, float u2) {
    float3 wh = fnormalize(v3(alpha * w.x, alpha * w.y, w.z));
    if (wh.z < 0.0f) wh = -wh;
    const float3 t1 = wh.z < 0.99999f ? fnormalize(v3(-wh.y, wh.x, 0.0f)) : v3(1.0f, 0.0f, 0.0f);
    const float3 t2 = cross(wh, t1);
    const float r = sqrtf(u1);
    float sn, cs;
    fsincos(2.0f * kPi * u2 - kPi, sn, cs);   // uniform angle in [-pi, pi): the accurate range of sin/cos.approx
    const float px = r * cs;
    float py = r * sn;
    const float h = sqrtf(fmaxf(0.0f, 1.0f - px * px));
    const float lerp_t = 0.5f + 0.5f * wh.z;
    py = (1.0f - lerp_t) * h + lerp_t * py;
    const float pz = sqrtf(fmaxf(0.0f, 1.0f - (px * px + py * py)));
    const float3 nh = px * t1 + py * t2 + pz * wh;
    return fnormalize(v3(alpha * nh.x, alpha * nh.y, fmaxf(1e-6f, nh.z)));
}
// unpolarised Fresnel reflectance of the coating (Dielectric.h:20-44)
B2PT_HD float fresnel(float cos_i) {
    cos_i = clampf(cos_i, -1.0f, 1.0f);
    float eta = kEta;
    if (cos_i < 0.0f) {
        eta = 1.0f / kEta;
        cos_i = -cos_i;
    }
    const float s2t = (1.0f - cos_i * cos_i) / (eta * eta);
    if (s2t >= 1.0f) return 1.0f;
    const float cos_t = sqrtf(1.0f - s2t);
    const float r_parl = (eta * cos_i - cos_t) / (eta * cos_i + cos_t);
    const float r_perp = (cos_i - eta * cos_t) / (cos_i + eta * cos_t);
    return 0.5f * (r_parl * r_parl + r_perp * r_perp);
}
// Snell refraction about n (Dielectric.h:68-92); false on total internal reflection
B2PT_HD bool refract_about(float3 wi, float3 n, float& etap, float3& wt) {
    float cos_i = dot(n, wi);
    float eta = kEta;
    if (cos_i < 0.0f) {
        eta = 1.0f / kEta;
        cos_i = -cos_i;
        n = -n;
    }
    const float s2t = fmaxf(0.0f, 1.0f - cos_i * cos_i) / (eta * eta);
    if (s2t >= 1.0f) return false;
    const float cos_t = sqrtf(1.0f - s2t);
    wt = -wi / eta + (cos_i / eta - cos_t) * n;
    etap = eta;
    return true;
}
B2PT_HD float transmittance(float3 w) {   // Tr over the slab thickness (GlossyDiffuse.h:97-106)
    return fexp(-kThickness / fabsf(w.z));
}

// ---- event at the top interface: Dielectric::Sample_f as a weight ---------------------------------
struct TopSample {
    float3 wi;
    float weight;      // f |cos_i| / pdf
    float pdf;         // only when asked for
    bool ok;
    bool reflected;
};
B2PT_HD void top_sample(float3 wo, float alpha, bool allow_r, bool allow_t, bool radiance, bool want_pdf, float uc,
                        float u1, float u2, TopSample& s) {
    s.ok = false;
    s.pdf = 0.0f;
    s.weight = 0.0f;
    s.reflected = false;
    s.wi = v3(0.0f, 0.0f, 1.0f);
    const bool smooth = alpha < 1e-3f;          // Surface.h:22
    const float a2 = alpha * alpha;
    float3 wm = v3(0.0f, 0.0f, 1.0f);
    if (!smooth) wm = sample_wm(wo, alpha, u1, u2);
    const float c = dot(wo, wm);
    const float R = fresnel(c), T = 1.0f - R;
    const float pr = allow_r ? R : 0.0f, pt = allow_t ? T : 0.0f;
    if (pr == 0.0f && pt == 0.0f) return;
    const float p_sum = pr + pt;
    float l1o = 1.0f, d = 1.0f;                 // 1 + Lambda(wo), D(wm)
    if (!smooth) {
        l1o = lambda1p(wo, a2);
        if (want_pdf) d = ggx_d(wm, a2);
    }
    // one tail for both lobes (a warp whose lanes sit in different walk phases runs it once): direction, the
    // hemisphere test of that lobe, G / G1, the lobe's extra factor
    const bool refl = uc < pr / p_sum;
    float etap = 1.0f;
    float3 wi = v3(0.0f, 0.0f, 1.0f);
    bool good;
    if (refl) {
        wi = smooth ? v3(-wo.x, -wo.y, wo.z) : -wo + (2.0f * c) * wm;
        good = wo.z * wi.z > 0.0f;
    } else {
        good = refract_about(wo, wm, etap, wi) && !(wo.z * wi.z > 0.0f) && wi.z != 0.0f;
    }
    if (!good) return;
    float g_ratio = 1.0f;                       // G(wo, wi) / G1(wo)
    if (!smooth) g_ratio = l1o / (l1o + lambda1p(wi, a2) - 1.0f);
    s.wi = wi;
    s.reflected = refl;
    s.weight = g_ratio * p_sum;
    if (!refl && radiance) s.weight /= etap * etap;
    if (want_pdf) {
        if (smooth) {
            s.pdf = (refl ? pr : pt) / p_sum;
        } else if (refl) {
            s.pdf = d / (l1o * 4.0f * fabsf(wo.z)) * pr / p_sum;
        } else {
            const float dwi = dot(wi, wm);
            const float sq = dwi + c / etap;
            s.pdf = d / (l1o * fabsf(wo.z)) * fabsf(c) * fabsf(dwi) / (sq * sq) * pt / p_sum;
        }
    }
    s.ok = s.weight > 0.0f && s.wi.z != 0.0f;   // usable(): f != 0, pdf != 0, wi.z != 0 (NaN fails the first test)
}

// Next-event term of a walk that sits at the base and looks up along `up` (z > 0) towards the light
// direction wi (z > 0) through the rough coating: Dielectric::f * MIS weight against the cosine pdf that
// produced `up` (GlossyDiffuse.h:337-355).  wdn = -up.  lam_wi = Lambda(wi).
B2PT_HD float exit_term(float3 up, float3 wi, float a2, float lam_wi) {
    const float3 wdn = -up;
    const float cos_o = wdn.z, cos_i = wi.z;
    if (cos_i == 0.0f || cos_o == 0.0f) return 0.0f;
    // transmission from inside (cos_o < 0) to outside (cos_i > 0): etap = 1 / eta (Dielectric.h:104-108)
    const float etap = 1.0f / kEta;
    float3 wm = wi * etap + wdn;
    const float l2 = dot(wm, wm);
    if (l2 == 0.0f) return 0.0f;
    wm = wm * frsqrt(l2);
    if (wm.z < 0.0f) wm = -wm;
    const float dwi = dot(wi, wm), dwo = dot(wdn, wm);
    if (dwi * cos_i < 0.0f || dwo * cos_o < 0.0f) return 0.0f;
    const float T = 1.0f - fresnel(dwo);
    if (!(T > 0.0f)) return 0.0f;
    const float sq = dwi + dwo * kEta;            // dwi + dwo / etap
    const float sq2 = sq * sq;
    const float d = ggx_d(wm, a2);
    const float l1o = lambda1p(wdn, a2);
    const float g = 1.0f / (l1o + lam_wi);
    const float f_exit = d * T * g * fabsf(dwi * dwo / (sq2 * cos_i * cos_o)) * (kEta * kEta);   // / etap^2
    const float exit_pdf = d / (l1o * fabsf(cos_o)) * fabsf(dwo) * fabsf(dwi) / sq2;
    const float bs_pdf = up.z * kInvPi;
    const float wt = bs_pdf * bs_pdf / (bs_pdf * bs_pdf + exit_pdf * exit_pdf);
    const float r = f_exit * wt;
    return r == r ? r : 0.0f;
}

// entrance reflection of `f` (GlossyDiffuse.h:185-187): Dielectric::f(wo, wi), both on the +z side
B2PT_HD float entrance_reflection(float3 wo, float3 wi, float alpha) {
    if (alpha < 1e-3f) return 0.0f;
    const float a2 = alpha * alpha;
    const float cos_o = wo.z, cos_i = wi.z;
    if (cos_i == 0.0f || cos_o == 0.0f) return 0.0f;
    float3 wm = wi + wo;
    const float l2 = dot(wm, wm);
    if (l2 == 0.0f) return 0.0f;
    wm = wm * frsqrt(l2);
    if (wm.z < 0.0f) wm = -wm;
    if (dot(wm, wi) * cos_i < 0.0f || dot(wm, wo) * cos_o < 0.0f) return 0.0f;
    const float F = fresnel(dot(wo, wm));
    const float g = 1.0f / (lambda1p(wo, a2) + lambda1p(wi, a2) - 1.0f);
    return ggx_d(wm, a2) * g * F / fabsf(4.0f * cos_i * cos_o);
}

// ---- a walk task -----------------------------------------------------------------------------------
struct Walk {
    // per hit
    float3 wo;         // local, flipped to z >= 0
    float3 wi;         // f walks: light direction, same flip
    float3 albedo;
    float alpha;
    float lam_wi;      // f walks over a rough coating: Lambda(wi)
    // state
    float3 w;          // direction of travel inside the slab
    float3 beta;       // throughput (sample walk: the weight so far)
    float3 f;          // f walk: the estimate so far
    float c_wis;       // f walk: the walk-constant factor of the next-event term along the pre-sampled wis
    int step;          // next step; < 0: finished
    bool is_f;
    bool ok;           // sample walk: a direction was produced
    bool flipped;      // wo.z was negative
};

// Lambda(wi) of an f walk's light direction: per hit, not per walk (the engine's prep kernel stores it)
B2PT_HD float walk_lambda_wi(float3 wi, float alpha) {
    return alpha < 1e-3f ? 0.0f : lambda1p(wi, alpha * alpha) - 1.0f;
}
B2PT_HD void walk_begin_f(Walk& k, float3 wo, float3 wi, float3 albedo, float alpha, float lam_wi) {
    k.flipped = wo.z < 0.0f;
    if (k.flipped) {   // two-sided (:153-157)
        wo = -wo;
        wi = -wi;
    }
    k.wo = wo;
    k.wi = wi;
    k.albedo = albedo;
    k.alpha = alpha;   // roughness^2, Surface.h:26
    k.lam_wi = lam_wi;
    k.w = v3(0.0f, 0.0f, -1.0f);
    k.beta = k.f = v3(0.0f);
    k.c_wis = 0.0f;
    k.step = 0;
    k.is_f = true;
    k.ok = false;
}
B2PT_HD void walk_begin_s(Walk& k, float3 wo, float3 albedo, float alpha) {
    k.flipped = wo.z < 0.0f;
    if (k.flipped) wo = -wo;
    k.wo = wo;
    k.wi = v3(0.0f, 0.0f, 1.0f);
    k.albedo = albedo;
    k.alpha = alpha;
    k.lam_wi = 0.0f;
    k.w = v3(0.0f, 0.0f, -1.0f);
    k.beta = k.f = v3(0.0f);
    k.c_wis = 0.0f;
    k.step = 0;
    k.is_f = false;
    k.ok = false;
}

// Russian roulette of both walk kinds (:230-247, :424-437): only past depth 3, only below 1/4
B2PT_HD bool walk_roulette(Walk& k, const WalkKey& key, int depth, int which) {
    const float m = max3(k.beta);
    bool survive = true;
    if (depth > 3 && m < 0.25f) {
        const float q = fmaxf(0.0f, 1.0f - m);
        survive = !(walk_rr_uniform(key, k.step, which) < q);
        k.beta = k.beta * (1.0f / (1.0f - q));
    }
    return survive;
}

// One step = [top event][bottom event].  Returns false when the task is finished:
//   f walk:      k.f is this walk's contribution to 5 f (the caller adds the entrance reflection and divides)
//   sample walk: k.ok, and on success k.w = sampled direction (un-flipped), k.beta = f |cos| / pdf
//
// Control flow is kept STRUCTURED on purpose — one `alive` flag, no return before the end: a warp of the engine
// holds walks of both kinds in different steps, and an early return inside a conditional moves the point where
// its lanes meet again to the end of the function, so that the two heavy blocks (the visible-normal sample of
// the top event, the next-event term of the bottom event) would run once per GROUP of lanes instead of once per
// warp (measured: 11.9 active lanes per instruction instead of 20, profiles/r02_k_layered_engine_*).
template <int RNG = 0>
B2PT_HD bool walk_step(Walk& k, const WalkKey& key) {
    const int st = k.step;
    const Draws u = walk_draws<RNG>(key, st);
    const bool smooth = k.alpha < 1e-3f;
    const float a2 = k.alpha * k.alpha;
    const bool is_f = k.is_f;
    const bool first = st == 0;
    const bool pre_b = is_f && st == 1;                 // the f walk's pre-sample of wis
    const bool walking = st >= (is_f ? 2 : 1);          // a top event of the walk proper

    // ---- top ----
    bool alive = true;
    if (walking) {
        alive = walk_roulette(k, key, is_f ? 2 * st - 3 : 2 * st - 1, 0) && k.w.z != 0.0f;
        k.beta *= transmittance(k.w);
    }
    const float3 dir = first ? k.wo : (pre_b ? k.wi : -k.w);
    TopSample ts;
    //            f walk: wos and wis transmit only, the walk reflects only; sample walk: both lobes
    top_sample(dir, k.alpha, /*allow_r=*/!is_f || walking, /*allow_t=*/!is_f || !walking, /*radiance=*/!pre_b,
               /*want_pdf=*/pre_b && !smooth, u.uc, u.u1, u.u2, ts);
    alive = alive && ts.ok;
    if (pre_b) {
        // next-event term along wis at every visit of the base (:318-334): everything but the walk's throughput and
        // the albedo is constant — diffuse f = albedo / pi, |wis.z|, MIS weight against the base's cosine pdf,
        // Tr(wis), wis.f / wis.pdf
        float wt = 1.0f;
        if (!smooth) {
            const float base_pdf = fabsf(ts.wi.z) * kInvPi;
            wt = ts.pdf * ts.pdf / (ts.pdf * ts.pdf + base_pdf * base_pdf);
        }
        k.c_wis = kInvPi * wt * transmittance(ts.wi) * ts.weight;
    } else {
        k.beta = first ? v3(ts.weight) : k.beta * ts.weight;
        k.w = ts.wi;
    }
    // the sample walk leaves the layers by reflecting off the top at once or by transmitting out from below;
    // its depth-9 event is the last chance (:513)
    const bool exits = alive && !is_f && (first ? ts.reflected : !ts.reflected);
    const bool last_s = !is_f && st == kStepsS - 1;
    // no bottom event in the first step of an f walk
    bool bottom = alive && !exits && !last_s && !(is_f && first);

    // ---- bottom ----
    if (bottom) {
        bottom = walk_roulette(k, key, is_f ? 2 * st - 2 : 2 * st, 1) && k.w.z != 0.0f;
        k.beta *= transmittance(k.w);
        if (bottom && is_f) k.f += k.beta * k.albedo * k.c_wis;
        // cosine sample of the base (LambertDiffuse.h:110-132): weight = albedo exactly
        const float r = sqrtf(u.d1);
        float sn, cs;
        fsincos(2.0f * kPi * u.d2 - kPi, sn, cs);
        const float3 up = v3(r * cs, r * sn, sqrtf(fmaxf(0.0f, 1.0f - u.d1)));
        k.beta *= k.albedo;
        bottom = bottom && max3(k.beta) > 0.0f && up.z != 0.0f;
        if (bottom && is_f && !smooth) k.f += k.beta * (transmittance(up) * exit_term(up, k.wi, a2, k.lam_wi));
        k.w = up;
    }

    // ---- what happens next ----
    if (exits) {
        if (k.flipped) k.w = -k.w;
        k.ok = true;
    }
    const bool f_first_ok = alive && is_f && first;
    const bool go_on = f_first_ok || (bottom && !(is_f && st == kStepsF - 1));
    k.step = go_on ? st + 1 : -1;
    return go_on;
}

}  // namespace fastlayer
}  // namespace b2pt
