// ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Compiles the reference's OWN device code for the host, straight from the
// sources where they lie under /root/reference (nothing is copied):
//   Renderer/OptiX/devicePrograms.cu   raygen / closest-hit / any-hit / miss
//   Renderer/OptiX/PBRT/*.h            Conductor, Dielectric, LambertDiffuse,
//                                      GlossyDiffuse, Microfacet, ...
// behind the stand-in OptiX headers of ref_shim/ (thread-local hit record,
// launch index, payload registers).  The result, _ref/liboracle_ref.so, is
//   * the parity oracle for BSDF f / Sample_f / PDF (ref_bsdf_*),
//   * the parity oracle for whole frames (ref_render: one
//     __raygen__renderFrame() per pixel per frame.id, OpenMP over rows), and
//   * the timed CPU baseline ("kind": "reference") of bench.py.
//
// What is NOT reference code here, and says so wherever a number appears:
//   * optixTrace: the reference calls the closed OptiX runtime
//     (devicePrograms.cu:224,248).  Here it is served by oracle/host_bvh.h with
//     the fp64 hit contract of oracle_trace.cpp.
//   * BSDF modes other than AUTO: the reference selects them by editing the
//     source (commented lines devicePrograms.cu:304-307, 324-327).  ref_mode_*
//     below restates __closesthit__radiance with exactly those lines switched
//     in; in AUTO mode the restatement is checked bit-for-bit against the
//     unmodified program (tests/test_bsdf_cpu.py::test_ref_oracle_reproduces_goldens).
//   * hasNormalTexture / hasMetalRoughTexture are set from their own texture
//     ids (the reference copies HasAlbedoTex() into all three,
//     OptixRenderer.cpp:535,540 — SURVEY Q4, documented FIX).
#include <omp.h>

#include <atomic>
#include <cstring>
#include <memory>
#include <vector>

#include "optix_device.h"  // ref_shim: must come before the reference sources

thread_local ShimThreadState g_shim;

#define __global__
#include "Renderer/OptiX/devicePrograms.cu"  // the reference, unmodified

#include "../include/b2pt.h"
#include "host_bvh.h"
#include "texunit.h"

// The reference declares `extern "C" __constant__ LaunchParams optixLaunchParams`
// (devicePrograms.cu:42); OptiX fills it per launch.  One copy per thread so
// that several frames can be rendered concurrently if a caller wishes.
extern "C" {
LaunchParams optixLaunchParams;
}

namespace {

struct RefScene {
    std::vector<MeshSBTData> sbt;                 // one record per mesh
    std::vector<std::vector<glm::vec2>> zero_uv;  // stand-in when a mesh has no texcoords
    std::vector<ShimTexture> textures;
    std::vector<float> soup;                      // world-space triangles, 9 floats each
    std::vector<int32_t> tri_mesh, tri_local;     // global prim -> mesh, local prim
    oracle::HostBvh bvh;
    // keep the caller's arrays alive
    std::vector<std::vector<float>> keep_f;
    std::vector<std::vector<int32_t>> keep_i;
    std::vector<std::vector<uint8_t>> keep_t;
};

thread_local const RefScene* g_scene = nullptr;
thread_local uint64_t g_radiance_rays = 0, g_shadow_rays = 0;
thread_local int g_mode = B2PT_BSDF_AUTO;
thread_local bool g_force_restated = false;  // AUTO through ref_closesthit_mode (self-check)

}  // namespace

// tex2D<float4> of the descriptor in OptixRenderer.cpp:592-603: the texture unit's filter as
// measured on a B200 (texunit.h).
float4 shim_tex2d(unsigned long long obj, float x, float y) {
    const ShimTexture& t = *reinterpret_cast<const ShimTexture*>(obj);
    float o[4];
    oracle::texunit_sample(t.rgba, t.width, t.height, x, y, o);
    return make_float4(o[0], o[1], o[2], o[3]);
}

static void ref_closesthit_mode();  // below

// ---------------------------------------------------------------------------
// optixTrace: NOT reference code (see header).  Hit contract = oracle_trace.cpp.
// ---------------------------------------------------------------------------
void optixTrace(OptixTraversableHandle, float3 origin, float3 direction, float tmin, float tmax, float,
                OptixVisibilityMask, unsigned int rayFlags, unsigned int sbtOffset, unsigned int,
                unsigned int, uint32_t& p0, uint32_t& p1) {
    const RefScene& sc = *g_scene;
    const ShimThreadState saved = g_shim;  // the shadow trace is nested inside closest-hit
    const float o[3] = {origin.x, origin.y, origin.z}, d[3] = {direction.x, direction.y, direction.z};
    const bool shadow = sbtOffset == SHADOW_RAY_TYPE;
    g_shim.p0 = p0;
    g_shim.p1 = p1;
    g_shim.ray_dir = direction;
    // any-hit program (devicePrograms.cu:549-561) as the acceptance filter
    auto accept = [&](int64_t prim, double u, double v) {
        const int m = sc.tri_mesh[prim];
        if (!sc.sbt[m].hasAlbedoTexture) return true;  // AlphaCutout returns false without a texture
        g_shim.sbt_data = (void*)&sc.sbt[m];
        g_shim.prim = (unsigned)sc.tri_local[prim];
        g_shim.bary = make_float2((float)u, (float)v);
        g_shim.ignored = false;
        if (shadow) __anyhit__shadow(); else __anyhit__radiance();
        return !g_shim.ignored;
    };
    if (shadow) {
        g_shadow_rays++;
        const bool occluded = sc.bvh.any(o, d, tmin, tmax, accept);
        (void)rayFlags;  // TERMINATE_ON_FIRST_HIT: closest-hit of the shadow type is empty (:56-59)
        if (!occluded) __miss__shadow();
    } else {
        g_radiance_rays++;
        const oracle::ClosestHit h = sc.bvh.closest(o, d, tmin, tmax, accept);
        if (h.prim < 0) {
            __miss__radiance();
        } else {
            const int m = sc.tri_mesh[h.prim];
            g_shim.sbt_data = (void*)&sc.sbt[m];
            g_shim.prim = (unsigned)sc.tri_local[h.prim];
            g_shim.bary = make_float2((float)h.u, (float)h.v);
            g_shim.backface = h.backface;
            if (g_mode == B2PT_BSDF_AUTO && !g_force_restated) __closesthit__radiance(); else ref_closesthit_mode();
        }
    }
    g_shim = saved;
}

// ---------------------------------------------------------------------------
// __closesthit__radiance (devicePrograms.cu:343-514) with the BSDF selected at
// run time.  Every helper called here is the reference's own function from the
// translation unit included above; only the two dispatchers GetNewRayDirection
// (:303-321) and BRDF (:323-341) are replaced by the alternatives the reference
// keeps as comments (:304-307, :324-327).  Debug printf blocks are dropped.
// ---------------------------------------------------------------------------
namespace {

bool mode_sample(int mode, unsigned int& seed, Surface& surface, BSDFSample& s) {
    switch (mode) {
        case B2PT_BSDF_CONDUCTOR:
            return PBRT::Conductor::Sample_f(seed, surface.albedo, surface.roughness, surface.outgoingRay, s);
        case B2PT_BSDF_DIFFUSE:
            return PBRT::LambertDiffuse::Sample_f(seed, surface, surface.outgoingRay, s, true);
        case B2PT_BSDF_DIELECTRIC:
            return PBRT::Dielectric::Sample_f(seed, surface, surface.outgoingRay, s,
                                              PBRT::Dielectric::TransportMode::Radiance, true, true);
        case B2PT_BSDF_LAYERED:
            return PBRT::GlossyDiffuse::Sample_f(seed, surface, s);
        default:
            return GetNewRayDirection(seed, surface, s);
    }
}

glm::vec3 mode_f(int mode, unsigned int& seed, Surface& surface, glm::vec3 wi) {
    switch (mode) {
        case B2PT_BSDF_CONDUCTOR:
            return PBRT::Conductor::f(surface.albedo, surface.roughness, surface.outgoingRay, wi);
        case B2PT_BSDF_DIFFUSE:
            return PBRT::LambertDiffuse::f(surface, surface.outgoingRay, wi);
        case B2PT_BSDF_DIELECTRIC:
            return PBRT::Dielectric::f(surface, surface.outgoingRay, wi, PBRT::Dielectric::TransportMode::Radiance);
        case B2PT_BSDF_LAYERED:
            return PBRT::GlossyDiffuse::f(seed, surface, wi);
        default:
            return BRDF(seed, surface, wi);
    }
}
}  // namespace

static void ref_closesthit_mode() {
    const glm::vec3 SHADING_SPACE_NORMAL = glm::vec3(0, 0, 1);
    MeshSBTData& sbtData = *(MeshSBTData*)optixGetSbtDataPointer();
    const bool isBackFaceHit = optixIsTriangleBackFaceHit();
    Surface surface;
    RadianceRayData* rayData = (RadianceRayData*)getPRD_1<RadianceRayData>();
    rayData->bounceCounter++;
    if (rayData->bounceCounter > rayData->maxBounces) {
        rayData->endPath = true;
        return;
    }
    surface.outgoingRay = glm::normalize(-OptixHelpers::Vec3(optixGetWorldRayDirection()));
    surface.index = sbtData.index[optixGetPrimitiveIndex()];
    GetVertices(sbtData, surface);
    GetNormal(sbtData, surface);
    if (isBackFaceHit) {
        surface.sNormal *= -1.0f;
        surface.gNormal *= -1.0f;
    }
    surface.position = GetSurfacePos(surface.vertices);
    surface.texCoord = GetTextureCoord(sbtData, surface.index);
    surface.albedo = sbtData.albedoColor;
    surface.metallic = sbtData.metallic;
    surface.roughness = sbtData.roughness;
    glm::vec3 normalTex = glm::vec3(0);
    SampleTextures(sbtData, normalTex, surface);
    surface.conductor = RandomOptix::rnd(rayData->randomSeed) < surface.metallic;
    if (normalTex != glm::vec3(0)) {
        glm::mat3 tbn = GetTBN(sbtData, surface);
        surface.sNormal = glm::normalize(tbn * (normalTex * glm::vec3(2.0f) - glm::vec3(1.0f)));
    }
    surface.ShadingToWorld = GetTBN(sbtData, surface);
    surface.WorldToShading = glm::transpose(surface.ShadingToWorld);
    surface.outgoingRay = surface.WorldToShading * surface.outgoingRay;

    float lightPropability = 0.0f;
    PointLight* pointLight = Lighting::GetRandomPointLight(&optixLaunchParams, rayData->randomSeed, lightPropability);
    if (lightPropability > 0.0f) {
        glm::vec3 lightDirection;
        const float lightVisibility = LightVisibility(pointLight->position, surface.position, surface.gNormal, lightDirection);
        lightDirection = surface.WorldToShading * lightDirection;
        if (lightVisibility > 0.0f) {
            glm::vec3 spectrum = mode_f(g_mode, rayData->randomSeed, surface, lightDirection) *
                                 AbsDot(lightDirection, SHADING_SPACE_NORMAL);
            if (spectrum != glm::vec3(0)) {
                rayData->radiance += (rayData->beta * spectrum * Lighting::GetSample(surface.position, pointLight)) /
                                     (lightPropability * Lighting::GetPDF(pointLight));
            }
        }
    }
    BSDFSample bsdfSample;
    if (!mode_sample(g_mode, rayData->randomSeed, surface, bsdfSample)) {
        rayData->endPath = true;
        return;
    }
    rayData->beta *= bsdfSample.color * AbsDot(bsdfSample.direction, SHADING_SPACE_NORMAL) / bsdfSample.pdf;
    glm::vec3 positionOffset = 1e-3f * surface.gNormal;
    if (glm::dot(bsdfSample.direction, SHADING_SPACE_NORMAL) < 0) positionOffset = -positionOffset;
    rayData->nextOrigin = surface.position + positionOffset;
    rayData->nextDirection = glm::normalize(surface.ShadingToWorld * bsdfSample.direction);
}

// ---------------------------------------------------------------------------
// scene hand-off: the same B2ptSceneDesc the product takes (include/b2pt.h),
// turned into what OptixRenderer::BuildSBT fills (OptixRenderer.cpp:520-554).
// ---------------------------------------------------------------------------
extern "C" {

void* ref_scene_create(const B2ptSceneDesc* desc) {
    auto sc = std::make_unique<RefScene>();
    sc->textures.resize((size_t)desc->n_textures);
    for (int i = 0; i < desc->n_textures; i++) {
        const B2ptTexture& t = desc->textures[i];
        sc->keep_t.emplace_back(t.rgba, t.rgba + (size_t)t.width * t.height * 4);
        sc->textures[i] = ShimTexture{sc->keep_t.back().data(), t.width, t.height};
    }
    sc->sbt.resize((size_t)desc->n_meshes);
    sc->zero_uv.resize((size_t)desc->n_meshes);
    for (int m = 0; m < desc->n_meshes; m++) {
        const B2ptMesh& me = desc->meshes[m];
        MeshSBTData& d = sc->sbt[m];
        memset(&d, 0, sizeof d);
        sc->keep_f.emplace_back(me.positions, me.positions + (size_t)me.n_vertices * 3);
        d.vertex = (glm::vec3*)sc->keep_f.back().data();
        if (me.normals) {
            sc->keep_f.emplace_back(me.normals, me.normals + (size_t)me.n_vertices * 3);
            d.normal = (glm::vec3*)sc->keep_f.back().data();
        }
        if (me.texcoords) {
            sc->keep_f.emplace_back(me.texcoords, me.texcoords + (size_t)me.n_vertices * 2);
            d.texcoord = (glm::vec2*)sc->keep_f.back().data();
        } else {  // GetTextureCoord (:131) reads texcoord unconditionally
            sc->zero_uv[m].assign((size_t)std::max(me.n_vertices, 1), glm::vec2(0));
            d.texcoord = sc->zero_uv[m].data();
        }
        sc->keep_i.emplace_back(me.indices, me.indices + (size_t)me.n_triangles * 3);
        d.index = (glm::ivec3*)sc->keep_i.back().data();
        memcpy(&d.modelMatrix, me.model, 64);  // column-major == glm
        d.albedoColor = glm::vec3(me.albedo[0], me.albedo[1], me.albedo[2]);
        d.roughness = me.roughness;
        d.metallic = me.metallic;
        auto tex = [&](int id) { return id >= 0 && id < desc->n_textures; };
        d.hasAlbedoTexture = tex(me.tex_albedo);
        d.hasNormalTexture = tex(me.tex_normal);
        d.hasMetalRoughTexture = tex(me.tex_metal_rough);
        if (d.hasAlbedoTexture) d.albedoTexture = (cudaTextureObject_t)&sc->textures[me.tex_albedo];
        if (d.hasNormalTexture) d.normalTexture = (cudaTextureObject_t)&sc->textures[me.tex_normal];
        if (d.hasMetalRoughTexture) d.metalRoughTexture = (cudaTextureObject_t)&sc->textures[me.tex_metal_rough];
        // AlphaCutout (:533) tests sbtData.texcoord: keep it null-equivalent when the mesh has none
        for (int t = 0; t < me.n_triangles; t++) {
            for (int k = 0; k < 3; k++) {
                const glm::vec3 v = d.vertex[me.indices[3 * t + k]];
                const glm::vec4 w = d.modelMatrix * glm::vec4(v, 1);  // GetVertices (:77-81)
                sc->soup.push_back(w.x);
                sc->soup.push_back(w.y);
                sc->soup.push_back(w.z);
            }
            sc->tri_mesh.push_back(m);
            sc->tri_local.push_back(t);
        }
    }
    sc->bvh.build(sc->soup.data(), (int64_t)sc->tri_mesh.size());
    return sc.release();
}

void ref_scene_destroy(void* h) { delete static_cast<RefScene*>(h); }

int64_t ref_scene_triangles(void* h, float* soup_out) {
    RefScene* sc = static_cast<RefScene*>(h);
    if (soup_out) memcpy(soup_out, sc->soup.data(), sc->soup.size() * 4);
    return (int64_t)sc->tri_mesh.size();
}

// Render frames frame_first .. frame_first + n_frames - 1 (frame.id values; the
// reference pre-increments to 1 before its first launch, OptixRenderer.cpp:623).
//   mean_out  (W*H*3, may be NULL): running mean exactly as AddPathtracedFrame.frag
//             :18-24 in "continuous" mode: fb = mix(fb, new, 1/samples)
//   frames_out(n_frames*W*H*3, may be NULL): every frame as Render() returns it
//   counters  (may be NULL): [0] radiance rays, [1] shadow rays, [2] NaN/Inf samples
//             (kept out of the mean; the GL path would poison the pixel and
//             WriteEXR zeroes it, WriteImage.cpp:52-55), [3] threads used
// mode = B2PT_BSDF_*; restated = 1 forces ref_closesthit_mode even in AUTO.
static void render_impl(void* h, int width, int height, int y0, int y1, int ystep, const float* cam_pos, const float* inv_view,
                        const float* inv_proj, const B2ptPointLight* lights, int n_lights, int max_bounces,
                        int mode, int restated, uint32_t frame_first, int n_frames, float* mean_out,
                        float* frames_out, uint64_t* counters, double* seconds) {
    RefScene* sc = static_cast<RefScene*>(h);
    std::vector<glm::vec3> color((size_t)width * height);
    std::vector<PointLight> pl((size_t)std::max(n_lights, 1));
    for (int i = 0; i < n_lights; i++) {
        pl[i].position = glm::vec3(lights[i].position[0], lights[i].position[1], lights[i].position[2]);
        pl[i].color = glm::vec3(lights[i].color[0], lights[i].color[1], lights[i].color[2]);
    }
    optixLaunchParams.frame.colorBuffer = color.data();
    optixLaunchParams.frame.size = glm::ivec2(width, height);
    optixLaunchParams.camera.position = glm::vec3(cam_pos[0], cam_pos[1], cam_pos[2]);
    memcpy(&optixLaunchParams.camera.inverseViewMatrix, inv_view, 64);
    memcpy(&optixLaunchParams.camera.inverseProjectionMatrix, inv_proj, 64);
    optixLaunchParams.pointlights = pl.data();
    optixLaunchParams.pointLightCount = n_lights;
    optixLaunchParams.maxBounces = max_bounces;
    optixLaunchParams.traversable = 0;
    uint64_t rr = 0, sr = 0, nans = 0;
    int threads = 1;
    if (mean_out) memset(mean_out, 0, (size_t)width * height * 12);
    const double t_begin = omp_get_wtime();
    for (int f = 0; f < n_frames; f++) {
        optixLaunchParams.frame.id = frame_first + (uint32_t)f;
#pragma omp parallel reduction(+ : rr, sr)
        {
            g_scene = sc;
            g_mode = mode;
            g_force_restated = restated != 0;
            g_radiance_rays = g_shadow_rays = 0;
#pragma omp single
            threads = omp_get_num_threads();
#pragma omp for schedule(dynamic, 1)
            for (int y = y0; y < y1; y += ystep)
                for (int x = 0; x < width; x++) {
                    g_shim.launch_index = make_uint3((unsigned)x, (unsigned)y, 0);
                    __raygen__renderFrame();
                }
            rr += g_radiance_rays;
            sr += g_shadow_rays;
        }
        if (frames_out) memcpy(frames_out + (size_t)f * width * height * 3, color.data(), (size_t)width * height * 12);
        if (mean_out) {
            const float w = 1.0f / (float)(f + 1);
            for (size_t p = 0; p < (size_t)width * height; p++) {
                glm::vec3 c = color[p];
                if (!(std::isfinite(c.x) && std::isfinite(c.y) && std::isfinite(c.z))) {
                    nans++;
                    c = glm::vec3(0);
                }
                for (int k = 0; k < 3; k++) {
                    float& fb = mean_out[3 * p + k];
                    fb = fb * (1.0f - w) + c[k] * w;  // glm/GLSL mix(x, y, a) = x*(1-a) + y*a
                }
            }
        }
    }
    if (seconds) *seconds = omp_get_wtime() - t_begin;
    if (counters) {
        counters[0] = rr;
        counters[1] = sr;
        counters[2] = nans;
        counters[3] = (uint64_t)threads;
    }
}

}  // extern "C"

extern "C" {
void ref_render(void* h, int width, int height, const float* cam_pos, const float* inv_view,
                const float* inv_proj, const B2ptPointLight* lights, int n_lights, int max_bounces,
                int mode, int restated, uint32_t frame_first, int n_frames, float* mean_out,
                float* frames_out, uint64_t* counters) {
    render_impl(h, width, height, 0, height, 1, cam_pos, inv_view, inv_proj, lights, n_lights, max_bounces, mode,
                restated, frame_first, n_frames, mean_out, frames_out, counters, nullptr);
}
// Rows y0, y0 + ystep, ... < y1 only (the timed CPU baseline of bench.py renders a
// bounded, evenly spread subset of the rows of the full-size frame);
// *seconds = wall time of the frame loop.
void ref_render_band(void* h, int width, int height, int y0, int y1, int ystep, const float* cam_pos, const float* inv_view,
                     const float* inv_proj, const B2ptPointLight* lights, int n_lights, int max_bounces,
                     int mode, uint32_t frame_first, int n_frames, uint64_t* counters, double* seconds) {
    render_impl(h, width, height, y0, y1, ystep < 1 ? 1 : ystep, cam_pos, inv_view, inv_proj, lights, n_lights, max_bounces, mode, 0,
                frame_first, n_frames, nullptr, nullptr, counters, seconds);
}
}

// ---------------------------------------------------------------------------
// BSDF entry points: the reference's functions, one call per row.
// bsdf = B2PT_BSDF_{CONDUCTOR,DIFFUSE,DIELECTRIC,LAYERED}; all vectors are in
// the local shading frame (z = normal).  seed[i] is the LCG state on entry,
// seed_out[i] the state on return (so draw counts are comparable too).
// flags: bit0 reflection, bit1 transmission, bit2 specular, bit3 glossy.
// ---------------------------------------------------------------------------
namespace {
thread_local RadianceRayData g_dummy_prd;

void arm_payload() {  // Conductor::Sample_f calls IsDebugRay() -> getPRD_1 (RayData.h:71-74)
    g_dummy_prd.isDebugRay = false;
    uint32_t a, b;
    packPointer(&g_dummy_prd, a, b);
    g_shim.p0 = a;
    g_shim.p1 = b;
}

Surface make_surface(const float* albedo, float roughness, const float* wo) {
    Surface s;
    memset(&s, 0, sizeof s);
    s.albedo = glm::vec3(albedo[0], albedo[1], albedo[2]);
    s.roughness = roughness;
    s.outgoingRay = glm::vec3(wo[0], wo[1], wo[2]);
    return s;
}
}  // namespace

extern "C" {

void ref_bsdf_f(int bsdf, int64_t n, const float* albedo, const float* roughness, const float* wo,
                const float* wi, const uint32_t* seed, float* f_out, uint32_t* seed_out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        arm_payload();
        Surface s = make_surface(albedo + 3 * i, roughness[i], wo + 3 * i);
        const glm::vec3 o = s.outgoingRay, in(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
        unsigned int sd = seed ? seed[i] : 0u;
        glm::vec3 f(0);
        switch (bsdf) {
            case B2PT_BSDF_CONDUCTOR: f = PBRT::Conductor::f(s.albedo, s.roughness, o, in); break;
            case B2PT_BSDF_DIFFUSE: f = PBRT::LambertDiffuse::f(s, o, in); break;
            case B2PT_BSDF_DIELECTRIC: f = PBRT::Dielectric::f(s, o, in, PBRT::Dielectric::TransportMode::Radiance); break;
            case B2PT_BSDF_LAYERED: f = PBRT::GlossyDiffuse::f(sd, s, in); break;
        }
        f_out[3 * i] = f.x; f_out[3 * i + 1] = f.y; f_out[3 * i + 2] = f.z;
        if (seed_out) seed_out[i] = sd;
    }
}

void ref_bsdf_sample(int bsdf, int64_t n, const float* albedo, const float* roughness, const float* wo,
                     const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                     uint8_t* flags_out, uint32_t* seed_out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        arm_payload();
        Surface s = make_surface(albedo + 3 * i, roughness[i], wo + 3 * i);
        unsigned int sd = seed[i];
        BSDFSample bs;
        memset(&bs, 0, sizeof bs);
        const bool good = mode_sample(bsdf, sd, s, bs);
        ok[i] = good;
        for (int k = 0; k < 3; k++) {
            f_out[3 * i + k] = good ? bs.color[k] : 0.0f;
            wi_out[3 * i + k] = good ? bs.direction[k] : 0.0f;
        }
        pdf_out[i] = good ? bs.pdf : 0.0f;
        flags_out[i] = good ? (uint8_t)((bs.reflection ? 1 : 0) | (bs.transmission ? 2 : 0) |
                                        (bs.specular ? 4 : 0) | (bs.glossy ? 8 : 0)) : 0;
        if (seed_out) seed_out[i] = sd;
    }
}

void ref_bsdf_pdf(int bsdf, int64_t n, const float* roughness, const float* wo, const float* wi,
                  int reflection, int transmission, float* pdf_out) {
    const float one[3] = {1, 1, 1};
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        Surface s = make_surface(one, roughness[i], wo + 3 * i);
        const glm::vec3 o = s.outgoingRay, in(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
        float p = 0.0f;
        if (bsdf == B2PT_BSDF_DIELECTRIC) p = PBRT::Dielectric::PDF(s, o, in, reflection != 0, transmission != 0);
        else if (bsdf == B2PT_BSDF_DIFFUSE) p = PBRT::LambertDiffuse::PDF(s, o, in, reflection != 0);
        pdf_out[i] = p;
    }
}

// Building blocks (A5, A18): exposed so unit-level parity can localise a mismatch.
uint32_t ref_tea16(uint32_t a, uint32_t b) { return RandomOptix::tea<16>(a, b); }
float ref_rnd(uint32_t* seed) { unsigned int s = *seed; float r = RandomOptix::rnd(s); *seed = s; return r; }
void ref_microfacet(int64_t n, const float* w, const float* wm, const float* alpha, float* d_out,
                    float* lambda_out, float* g1_out, float* pdf_out) {
    for (int64_t i = 0; i < n; i++) {
        const glm::vec3 a(w[3 * i], w[3 * i + 1], w[3 * i + 2]), m(wm[3 * i], wm[3 * i + 1], wm[3 * i + 2]);
        d_out[i] = PBRT::Microfacet::D_Isotropic(m, alpha[i]);
        lambda_out[i] = PBRT::Microfacet::Lambda_Anisotropic(a, glm::vec2(alpha[i]));
        g1_out[i] = PBRT::Microfacet::G1_Anisotropic(a, glm::vec2(alpha[i]));
        pdf_out[i] = PBRT::Microfacet::PDF_Isotropic(a, m, alpha[i]);
    }
}
void ref_sample_wm(int64_t n, const float* w, const float* alpha, const uint32_t* seed, float* wm_out,
                   uint32_t* seed_out) {
    for (int64_t i = 0; i < n; i++) {
        unsigned int s = seed[i];
        const glm::vec3 m = PBRT::Microfacet::Sample_wm(s, glm::vec3(w[3 * i], w[3 * i + 1], w[3 * i + 2]), alpha[i]);
        wm_out[3 * i] = m.x; wm_out[3 * i + 1] = m.y; wm_out[3 * i + 2] = m.z;
        if (seed_out) seed_out[i] = s;
    }
}
void ref_fresnel(int64_t n, const float* cos_i, const float* reflectance, float* fr_dielectric, float* fr_complex) {
    for (int64_t i = 0; i < n; i++) {
        fr_dielectric[i] = PBRT::Dielectric::FresnelDielectric(cos_i[i], 1.5f);
        const glm::vec3 f = PBRT::Conductor::FresnelComplex(
            cos_i[i], glm::vec3(reflectance[3 * i], reflectance[3 * i + 1], reflectance[3 * i + 2]));
        fr_complex[3 * i] = f.x; fr_complex[3 * i + 1] = f.y; fr_complex[3 * i + 2] = f.z;
    }
}
// Camera() (devicePrograms.cu:601-623) for every pixel: W*H*3 directions.
void ref_camera_rays(int width, int height, const float* cam_pos, const float* inv_view, const float* inv_proj,
                     float* org_out, float* dir_out) {
    optixLaunchParams.frame.size = glm::ivec2(width, height);
    optixLaunchParams.camera.position = glm::vec3(cam_pos[0], cam_pos[1], cam_pos[2]);
    memcpy(&optixLaunchParams.camera.inverseViewMatrix, inv_view, 64);
    memcpy(&optixLaunchParams.camera.inverseProjectionMatrix, inv_proj, 64);
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            glm::vec3 o;
            glm::vec4 d;
            Camera(glm::ivec2(x, y), o, d);
            const size_t p = (size_t)y * width + x;
            org_out[3 * p] = o.x; org_out[3 * p + 1] = o.y; org_out[3 * p + 2] = o.z;
            dir_out[3 * p] = d.x; dir_out[3 * p + 1] = d.y; dir_out[3 * p + 2] = d.z;
        }
}

}  // extern "C"
