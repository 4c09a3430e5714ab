// integrator.cuh — per-thread logic of the wavefront stages (K4, K6, K7).
//
// The reference runs one path per thread inside __raygen__renderFrame
// (devicePrograms.cu:625-706) and shades inside __closesthit__radiance
// (:343-514).  Here the same arithmetic is cut at the two places where the
// reference calls optixTrace, giving three pure functions of plain structs:
//
//   camera_ray      Camera()                  :601-623   -> K4
//   surface_stage   closest-hit part 1        :366-452   -> K6  (ends by emitting
//                   the shadow ray of LightVisibility :216-241)
//   shade_stage     closest-hit part 2        :453-509   -> K7a-d (NEE add, BSDF
//                   sample, throughput, next ray, loop test of :646)
//
// They are B2PT_HD: the kernels in wavefront.cu call them after loading the SoA
// state, and the CPU test tier (tests/hostsim) runs them in a per-pixel loop to
// compare whole frames bit-for-bit with the host-compiled reference.
#pragma once

#include "bsdf.cuh"
#include "scene.cuh"

namespace b2pt {

struct CameraParams {   // LaunchParams.camera + frame.size (LaunchParams.h:11-21)
    float3 position;
    float inv_view[16];
    float inv_proj[16];
    int32_t width, height;
};

// glm's 4-component dot: (x*x + y*y) + (z*z + w*w)  (glm/detail/func_geometric.inl)
B2PT_HD float4 normalize4(float4 v) {
    const float d = (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    const float inv = 1.0f / sqrtf(d);
    return make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv);
}

// Pixel centre, no jitter.
B2PT_HD void camera_ray(const CameraParams& cam, int x, int y, float3& origin, float3& dir) {
    const float xs = ((float)x + 0.5f) / (float)cam.width;
    const float ys = ((float)y + 0.5f) / (float)cam.height;
    const float xn = xs * 2.0f - 1.0f, yn = ys * 2.0f - 1.0f;
    const float4 pv = mat4_mul(cam.inv_proj, xn, yn, 1.0f, 1.0f);
    const float4 pw = mat4_mul(cam.inv_view, pv.x, pv.y, pv.z, 0.0f);
    const float4 n = normalize4(pw);
    dir = v3(n.x, n.y, n.z);
    origin = cam.position;
}

// ---- RNG plumbing: how a stage opens/closes its generator -------------------
struct RngCtx {
    uint64_t seed;       // Philox key (B2ptConfig.seed)
    uint32_t pixel;      // y*W + x
    uint32_t sample;     // absolute sample index = frame.id of this sample
    uint32_t bounce;     // 0-based
    uint32_t lcg_state;  // RadianceRayData.randomSeed (LCG policies)
};
template <class Rng>
struct RngIO;
template <bool H>
struct RngIO<LcgRng<H>> {
    static B2PT_HD LcgRng<H> open(const RngCtx& c, uint32_t) { return LcgRng<H>(c.lcg_state); }
    static B2PT_HD uint32_t close(const LcgRng<H>& r) { return r.s; }
    // SamplePath seeds with tea<16>(W*y + x, frame.id) (:631)
    static B2PT_HD uint32_t first_state(uint32_t pixel, uint32_t frame_id) { return tea<16>(pixel, frame_id); }
};
template <>
struct RngIO<PhiloxRng> {
    static B2PT_HD PhiloxRng open(const RngCtx& c, uint32_t stage) {
        return PhiloxRng(c.seed, c.pixel, c.sample, c.bounce | (stage << 16));
    }
    static B2PT_HD uint32_t close(const PhiloxRng&) { return 0u; }
    static B2PT_HD uint32_t first_state(uint32_t, uint32_t) { return 0u; }
};

// ---- textures ---------------------------------------------------------------
#if defined(__CUDACC__)
struct DeviceTexFetch {
    const cudaTextureObject_t* objs;
    __device__ __forceinline__ float4 operator()(int id, float x, float y) const {
        return tex2D<float4>(objs[id], x, y);
    }
};
#endif

B2PT_HD float srgb8_to_linear_1(float c) {   // SRGB8ToLinear, one channel (:62-73)
    const float m = c < 0.04045f ? 0.0f : 1.0f;   // glm::step(edge, x)
    const float a = c / 12.92f;
#if defined(__CUDA_ARCH__)
    // the host library's powf is correctly rounded on all but 0.1% of the 65536 values a 16-bit texture
    // read can take (measured); CUDA's powf is a 4-ulp function, so round a double-precision pow instead
    const float b = (float)pow((double)((c + 0.055f) / 1.055f), (double)2.4f);
#else
    const float b = powf((c + 0.055f) / 1.055f, 2.4f);
#endif
    return a * (1.0f - m) + b * m;
}

// ---- K6: hit -> surface ------------------------------------------------------
struct SurfacePoint {
    float3 p, ng, ns;      // surface.position, gNormal, sNormal (final, after normal mapping)
    ShadePoint sp;         // albedo, roughness (after textures)
    bool conductor;        // surface.conductor
    int light;             // chosen light or -1 (no lights)
    float3 sh_o, sh_d;     // shadow ray
    float sh_tmax;
};

B2PT_HD void tangent_frame(float3 n, float3& t, float3& b) {   // BuildTangentSpace (:168-184)
    const float3 c1 = cross(n, v3(0.0f, 0.0f, 1.0f));
    const float3 c2 = cross(n, v3(0.0f, 1.0f, 0.0f));
    t = length(c1) > length(c2) ? c1 : c2;
    t = normalize(t);
    b = cross(t, n);
}
B2PT_HD float3 to_local(float3 t, float3 b, float3 n, float3 v) {   // transpose(mat3(T,B,N)) * v
    return v3(dot(t, v), dot(b, v), dot(n, v));
}
B2PT_HD float3 to_world(float3 t, float3 b, float3 n, float3 v) {   // mat3(T,B,N) * v
    return v3(t.x * v.x + b.x * v.y + n.x * v.z, t.y * v.x + b.y * v.y + n.y * v.z,
              t.z * v.x + b.z * v.y + n.z * v.z);
}

// `mode` = B2PT_BSDF_*; in a forced mode the class draw is still consumed (Q8).
// Meshes without vertex normals shade with Ns = Ng (documented FIX: the
// reference would normalise a zero vector, devicePrograms.cu:94-113).
template <class Rng, class TexFetch>
B2PT_HD void surface_stage(const DeviceScene& sc, const TexFetch& tex, const HitRecord& hit, float3 ray_d,
                           const B2ptPointLight* lights, int n_lights, Rng& rng, SurfacePoint& out) {
    const uint32_t slot = hit.tri & ~kHitBackface;
    const bool backface = (hit.tri & kHitBackface) != 0;
    const float u = hit.u, v = hit.v;
    const TriRecord* tp = sc.tris + slot;
    const float4 ta = B2PT_LDG4(&tp->a), tb = B2PT_LDG4(&tp->b), tc = B2PT_LDG4(&tp->c);
    const float3 v0 = v3(ta.x, ta.y, ta.z), v1 = v3(ta.w, tb.x, tb.y), v2 = v3(tb.z, tb.w, tc.x);
    const uint32_t mesh = f2u_bits(tc.z);
    const MaterialRecord& mat = sc.materials[mesh];

    const float3 wo = normalize(-ray_d);
    float3 ng = cross(v1 - v0, v2 - v0);
    const NrmRecord* np = sc.nrms + slot;
    const float4 na = B2PT_LDG4(&np->a), nb = B2PT_LDG4(&np->b), nc = B2PT_LDG4(&np->c);
    const bool has_normals = f2u_bits(nc.y) != 0u;
    const float w0 = 1.0f - u - v;
    float3 ns = v3(0.0f);
    if (has_normals) {
        const float3 n0 = v3(na.x, na.y, na.z), n1 = v3(na.w, nb.x, nb.y), n2 = v3(nb.z, nb.w, nc.x);
        const float3 no = w0 * n0 + u * n1 + v * n2;
        const float* m = mat.mcol;   // modelMatrix * vec4(Ns, 0), then the vec4 normalize (:99)
        const float4 nw = make_float4(B2PT_FADD(B2PT_FADD(B2PT_FMUL(m[0], no.x), B2PT_FMUL(m[4], no.y)), B2PT_FMUL(m[8], no.z)),
                                      B2PT_FADD(B2PT_FADD(B2PT_FMUL(m[1], no.x), B2PT_FMUL(m[5], no.y)), B2PT_FMUL(m[9], no.z)),
                                      B2PT_FADD(B2PT_FADD(B2PT_FMUL(m[2], no.x), B2PT_FMUL(m[6], no.y)), B2PT_FMUL(m[10], no.z)),
                                      B2PT_FADD(B2PT_FADD(B2PT_FMUL(m[3], no.x), B2PT_FMUL(m[7], no.y)), B2PT_FMUL(m[11], no.z)));
        const float4 nn = normalize4(nw);
        ns = v3(nn.x, nn.y, nn.z);
    }
    if (dot(wo, ng) < 0.0f) ng = -ng;
    ng = normalize(ng);
    if (!has_normals) ns = ng;
    if (dot(ng, ns) < 0.0f) ns = -ns;
    ns = normalize(ns);
    if (backface) {   // "normals always point outside" (:379-382)
        ns = ns * -1.0f;
        ng = ng * -1.0f;
    }
    out.p = w0 * v0 + u * v1 + v * v2;

    float3 albedo = v3(mat.albedo[0], mat.albedo[1], mat.albedo[2]);
    float metallic = mat.metallic, roughness = mat.roughness;
    float3 normal_tex = v3(0.0f);
    if (mat.tex_albedo >= 0 || mat.tex_normal >= 0 || mat.tex_metal_rough >= 0) {
        float tx = 0.0f, ty = 0.0f;
        if (sc.uvs) {
            const UvRecord ur = sc.uvs[slot];
            tx = w0 * ur.t0.x + u * ur.t1.x + v * ur.t2.x;
            ty = w0 * ur.t0.y + u * ur.t1.y + v * ur.t2.y;
        }
        if (mat.tex_albedo >= 0) {   // SampleTextures (:143-166)
            const float4 t = tex(mat.tex_albedo, tx, ty);
            albedo *= v3(srgb8_to_linear_1(t.x), srgb8_to_linear_1(t.y), srgb8_to_linear_1(t.z));
        }
        if (mat.tex_normal >= 0) {
            const float4 t = tex(mat.tex_normal, tx, ty);
            normal_tex = v3(t.x, t.y, t.z);
        }
        if (mat.tex_metal_rough >= 0) {
            const float4 t = tex(mat.tex_metal_rough, tx, ty);
            metallic = t.x;
            roughness = t.y;
        }
    }
    out.conductor = rng.next() < metallic;   // always drawn (:400)
    if (nonzero(normal_tex)) {   // (:403-409)
        float3 t, b;
        tangent_frame(ns, t, b);
        ns = normalize(to_world(t, b, ns, normal_tex * v3(2.0f) - v3(1.0f)));
    }
    out.ng = ng;
    out.ns = ns;
    out.sp.albedo = albedo;
    out.sp.roughness = roughness;

    // Lighting::GetRandomPointLight (LightMethods.h:25-41): a draw only if count > 1
    out.light = -1;
    out.sh_tmax = -1.0f;
    out.sh_o = out.sh_d = v3(0.0f);
    if (n_lights == 1) {
        out.light = 0;
    } else if (n_lights > 1) {
        const float r = rng.next();
        out.light = (int)(r * (float)n_lights);
    }
    if (out.light >= 0) {   // LightVisibility (:216-241): un-offset direction, full length (Q6)
        const B2ptPointLight& L = lights[out.light];
        const float3 ld = v3(L.position[0], L.position[1], L.position[2]) - out.p;
        out.sh_o = out.p + 1e-3f * ng;
        out.sh_d = normalize(ld);
        out.sh_tmax = length(ld);
    }
}

// ---- K7: shade ----------------------------------------------------------------
struct PathVertex {      // the part of RadianceRayData (RayData.h:5-20) a bounce updates
    float3 beta, radiance;
    float3 next_o, next_d;
};

// Returns true when the path continues (the loop test of SamplePath :646 for
// the NEXT iteration is applied by the caller, which knows bounce/max_bounces).
// `visible` = LightVisibility(...) > 0 for the shadow ray emitted by
// surface_stage; ray_d = direction of the ray that produced the hit.
template <int KIND, class Rng>
B2PT_HD bool shade_stage(const SurfacePoint& s, bool visible, float3 ray_d, const B2ptPointLight* lights,
                         int n_lights, Rng& rng, PathVertex& pv) {
    float3 t, b;
    tangent_frame(s.ns, t, b);   // GetTBN on the final shading normal (:412-413)
    const float3 n = s.ns;
    const float3 wo = to_local(t, b, n, normalize(-ray_d));
    const float3 up = v3(0.0f, 0.0f, 1.0f);

    if (s.light >= 0) {   // lightPropability > 0 (:449)
        const B2ptPointLight& L = lights[s.light];
        const float3 lpos = v3(L.position[0], L.position[1], L.position[2]);
        if (visible) {
            const float3 wl = to_local(t, b, n, normalize(lpos - s.p));
            const float3 spectrum = bsdf_f<KIND>(rng, s.sp, wo, wl) * absdot(wl, up);
            if (nonzero(spectrum)) {
                const float p_light = n_lights == 1 ? 1.0f : 1.0f / (float)n_lights;
                const float3 dl = s.p - lpos;   // Lighting::DistanceSquared(light, surface) (LightMethods.h:13-23)
                const float d2 = dl.x * dl.x + dl.y * dl.y + dl.z * dl.z;
                const float3 li = v3(L.color[0], L.color[1], L.color[2]) / d2;
                pv.radiance += (pv.beta * spectrum * li) / (p_light * 1.0f);
            }
        }
    }
    BsdfSample bs;
    if (!bsdf_sample<KIND>(rng, s.sp, wo, bs)) return false;   // endPath (:474-483)
    pv.beta *= bs.f * absdot(bs.wi, up) / bs.pdf;
    float3 offset = 1e-3f * s.ng;
    if (dot(bs.wi, up) < 0.0f) offset = -offset;
    pv.next_o = s.p + offset;
    pv.next_d = normalize(to_world(t, b, n, bs.wi));
    return true;
}

// loop test of SamplePath (:646) before tracing bounce `bounce` (0-based)
B2PT_HD bool path_continues(float3 beta, int bounce, int max_bounces) {
    return bounce < max_bounces && length(beta) > 0.00001f;
}

}  // namespace b2pt
