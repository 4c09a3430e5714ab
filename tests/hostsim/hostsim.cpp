// hostsim.cpp — TEST HARNESS: runs the per-thread logic of the CUDA kernels
// (the B2PT_HD functions in optixpathtracer_b200/csrc/*.cuh) on the host so
// that traversal / BVH-topology / BSDF logic can be checked against the oracle
// in the CPU-only test tier.  It is never part of the product: libb2pt.so does
// not contain it and nothing in the package loads it.
//
// The LBVH here is built by a sequential loop over the SAME per-element
// functions the build kernels call (morton63, karras_node, aabb_*), so a
// topology bug shows up without a GPU.
#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../optixpathtracer_b200/csrc/traverse.cuh"

using namespace b2pt;

namespace {

struct HostBvh {
    std::vector<BvhNode2> nodes;
    std::vector<TriRecord> tris;
    std::vector<uint32_t> order;
};

Aabb box_of(const std::vector<int2>& children, const std::vector<Aabb>& node_box,
            const std::vector<Aabb>& leaf_box, int c) {
    return c < 0 ? leaf_box[~c] : node_box[c];
}

void build(const float* soup, int n, HostBvh& out) {
    out.nodes.clear();
    out.tris.clear();
    out.order.clear();
    if (n == 0) return;
    std::vector<float3> wv(3 * (size_t)n);
    memcpy(wv.data(), soup, sizeof(float) * 9 * (size_t)n);
    float3 clo = v3(INFINITY), chi = v3(-INFINITY);
    float maxabs = 0.f;
    for (int g = 0; g < n; g++) {
        Aabb b = aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]);
        float3 c = (b.lo + b.hi) * 0.5f;
        clo = v3(fminf(clo.x, c.x), fminf(clo.y, c.y), fminf(clo.z, c.z));
        chi = v3(fmaxf(chi.x, c.x), fmaxf(chi.y, c.y), fmaxf(chi.z, c.z));
        const float vals[6] = {b.lo.x, b.lo.y, b.lo.z, b.hi.x, b.hi.y, b.hi.z};
        for (float f : vals) maxabs = fmaxf(maxabs, fabsf(f));
    }
    const float pad = fmaxf(maxabs, 1e-30f) * 3.8146973e-6f;
    const float3 ext = chi - clo;
    const float3 inv = v3(ext.x > 0 ? 1.0f / ext.x : 0.f, ext.y > 0 ? 1.0f / ext.y : 0.f, ext.z > 0 ? 1.0f / ext.z : 0.f);
    std::vector<uint64_t> keys(n);
    for (int g = 0; g < n; g++) {
        Aabb b = aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]);
        keys[g] = morton63((b.lo + b.hi) * 0.5f, clo, inv);
    }
    out.order.resize(n);
    std::iota(out.order.begin(), out.order.end(), 0u);
    std::stable_sort(out.order.begin(), out.order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> sk(n);
    for (int s = 0; s < n; s++) sk[s] = keys[out.order[s]];
    std::vector<Aabb> leaf_box(n);
    out.tris.resize(n);
    for (int s = 0; s < n; s++) {
        const uint32_t g = out.order[s];
        leaf_box[s] = aabb_pad(aabb_of_tri(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2]), pad);
        out.tris[s] = tri_pack(wv[3 * g], wv[3 * g + 1], wv[3 * g + 2], g, 0);
    }
    if (n == 1) {
        BvhNode2 nd;
        const Aabb b = leaf_box[0];
        nd.c0xy = make_float4(b.lo.x, b.hi.x, b.lo.y, b.hi.y);
        nd.c1xy = make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);
        nd.cz = make_float4(b.lo.z, b.hi.z, INFINITY, -INFINITY);
        nd.child0 = ~0; nd.child1 = ~0; nd.pad0 = nd.pad1 = 0;
        out.nodes.push_back(nd);
        return;
    }
    std::vector<int2> children(n - 1);
    std::vector<int> first(n - 1), last(n - 1);
    for (int i = 0; i < n - 1; i++) {
        int f, l, sp;
        karras_node(sk.data(), n, i, f, l, sp);
        first[i] = f; last[i] = l;
        children[i] = make_int2(sp == f ? ~sp : sp, sp + 1 == l ? ~(sp + 1) : sp + 1);
    }
    // refit: process nodes by decreasing range length is not needed — recurse.
    std::vector<Aabb> node_box(n - 1);
    std::vector<char> done(n - 1, 0);
    std::vector<int> stack;
    stack.push_back(0);
    while (!stack.empty()) {
        const int i = stack.back();
        const int2 ch = children[i];
        bool ready = true;
        if (ch.x >= 0 && !done[ch.x]) { stack.push_back(ch.x); ready = false; }
        if (ch.y >= 0 && !done[ch.y]) { stack.push_back(ch.y); ready = false; }
        if (!ready) continue;
        node_box[i] = aabb_union(box_of(children, node_box, leaf_box, ch.x), box_of(children, node_box, leaf_box, ch.y));
        done[i] = 1;
        stack.pop_back();
    }
    out.nodes.resize(n - 1);
    for (int i = 0; i < n - 1; i++) {
        const int2 ch = children[i];
        const Aabb b0 = box_of(children, node_box, leaf_box, ch.x), b1 = box_of(children, node_box, leaf_box, ch.y);
        BvhNode2 nd;
        nd.c0xy = make_float4(b0.lo.x, b0.hi.x, b0.lo.y, b0.hi.y);
        nd.c1xy = make_float4(b1.lo.x, b1.hi.x, b1.lo.y, b1.hi.y);
        nd.cz = make_float4(b0.lo.z, b0.hi.z, b1.lo.z, b1.hi.z);
        nd.child0 = ch.x; nd.child1 = ch.y; nd.pad0 = nd.pad1 = 0;
        out.nodes[i] = nd;
    }
}

HostBvh g_bvh;

}  // namespace

extern "C" {

// Build the LBVH over a world-space triangle soup (nt*9 floats). Returns the node count.
int hostsim_build(const float* soup, int nt) {
    build(soup, nt, g_bvh);
    return (int)g_bvh.nodes.size();
}

// Structural self-check: every leaf slot reachable exactly once, child boxes
// contain their triangles. Returns 0 when consistent.
int hostsim_check_topology() {
    const int n = (int)g_bvh.tris.size();
    if (n <= 1) return 0;
    std::vector<int> seen(n, 0);
    std::vector<int> st{0};
    int visited = 0;
    while (!st.empty()) {
        int c = st.back(); st.pop_back();
        if (c < 0) { if (~c >= n) return 2; seen[~c]++; continue; }
        if (c >= n - 1) return 3;
        visited++;
        st.push_back(g_bvh.nodes[c].child0);
        st.push_back(g_bvh.nodes[c].child1);
    }
    if (visited != n - 1) return 4;
    for (int s = 0; s < n; s++) if (seen[s] != 1) return 5;
    return 0;
}

void hostsim_trace_closest(const float* org, const float* dir, int64_t n, float tmin, float tmax,
                           int32_t* prim, float* t, float* u, float* v, uint8_t* backface,
                           uint64_t* counters /* nodes, tris, exact */) {
    uint64_t cn = 0, ct = 0, ce = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : cn, ct, ce)
    for (int64_t i = 0; i < n; i++) {
        HitRecord h;
        TraceStats ts = {0, 0, 0};
        traverse<false, true>(g_bvh.nodes.data(), g_bvh.tris.data(), (int)g_bvh.nodes.size(),
                              v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]),
                              v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), tmin, tmax, h, &ts);
        const bool hit = h.tri != kHitMiss;
        prim[i] = hit ? (int32_t)g_bvh.order[h.tri & ~kHitBackface] : -1;
        t[i] = hit ? h.t : 0.f; u[i] = hit ? h.u : 0.f; v[i] = hit ? h.v : 0.f;
        if (backface) backface[i] = hit && (h.tri & kHitBackface);
        cn += ts.nodes; ct += ts.tris; ce += ts.exact;
    }
    if (counters) { counters[0] = cn; counters[1] = ct; counters[2] = ce; }
}

void hostsim_trace_any(const float* org, const float* dir, int64_t n, float tmin, const float* tmax,
                       uint8_t* occluded) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        HitRecord h;
        occluded[i] = traverse<true, false>(g_bvh.nodes.data(), g_bvh.tris.data(), (int)g_bvh.nodes.size(),
                                            v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]),
                                            v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), tmin, tmax[i], h,
                                            nullptr) ? 1 : 0;
    }
}

}  // extern "C"

// ---- BSDF rows (bsdf_rows.cuh) on the host ---------------------------------
#include "../../include/b2pt.h"
#include "../../optixpathtracer_b200/csrc/bsdf_rows.cuh"

extern "C" {

void hostsim_bsdf_f(int bsdf, int lcg_mode, int64_t n, const float* albedo, const float* roughness,
                    const float* wo, const float* wi, const uint32_t* seed, float* f_out, uint32_t* seed_out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        if (lcg_mode == B2PT_RNG_PHILOX && bsdf == B2PT_BSDF_LAYERED) {   // the production walk engine
            fast_layered_row_f(i, albedo, roughness, wo, wi, seed, f_out);
            if (seed_out) seed_out[i] = 0u;
        } else if (lcg_mode == B2PT_RNG_PHILOX)
            bsdf_row_f<PhiloxRng>(bsdf, i, albedo, roughness, wo, wi, seed, f_out, seed_out);
        else if (lcg_mode == B2PT_RNG_LCG_HOST)
            bsdf_row_f<LcgRng<true>>(bsdf, i, albedo, roughness, wo, wi, seed, f_out, seed_out);
        else
            bsdf_row_f<LcgRng<false>>(bsdf, i, albedo, roughness, wo, wi, seed, f_out, seed_out);
    }
}

void hostsim_bsdf_sample(int bsdf, int lcg_mode, int64_t n, const float* albedo, const float* roughness,
                         const float* wo, const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out,
                         float* pdf_out, uint8_t* flags_out, uint32_t* seed_out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        if (lcg_mode == B2PT_RNG_PHILOX && bsdf == B2PT_BSDF_LAYERED) {
            fast_layered_row_sample(i, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out, flags_out);
            if (seed_out) seed_out[i] = 0u;
        } else if (lcg_mode == B2PT_RNG_PHILOX)
            bsdf_row_sample<PhiloxRng>(bsdf, i, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out, flags_out, seed_out);
        else if (lcg_mode == B2PT_RNG_LCG_HOST)
            bsdf_row_sample<LcgRng<true>>(bsdf, i, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out,
                                          flags_out, seed_out);
        else
            bsdf_row_sample<LcgRng<false>>(bsdf, i, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out,
                                           flags_out, seed_out);
    }
}

// kind 0 sin, 1 cos, 2 exp: hostmath.cuh's restatement beside this host's libm
void hostsim_hostmath(int kind, int64_t n, const float* x, float* restated, float* libm) {
    for (int64_t i = 0; i < n; i++) {
        restated[i] = kind == 0 ? hostmath::sinf_host(x[i]) : kind == 1 ? hostmath::cosf_host(x[i]) : hostmath::expf_host(x[i]);
        libm[i] = kind == 0 ? ::sinf(x[i]) : kind == 1 ? ::cosf(x[i]) : ::expf(x[i]);
    }
}

void hostsim_bsdf_pdf(int bsdf, int64_t n, const float* roughness, const float* wo, const float* wi,
                      int reflection, int transmission, float* pdf_out) {
    for (int64_t i = 0; i < n; i++)
        bsdf_row_pdf(bsdf, i, roughness, wo, wi, reflection != 0, transmission != 0, pdf_out);
}

}  // extern "C"

// ---- whole frames: the wavefront stages (integrator.cuh) run per pixel --------
#include "../../optixpathtracer_b200/csrc/alpha.cuh"
#include "../../oracle/texunit.h"

namespace {

struct HostTexFetch {   // the texture unit as measured (oracle/texunit.h), as in oracle/ref_harness.cpp
    const B2ptTexture* tex;
    float4 operator()(int id, float x, float y) const {
        const B2ptTexture& t = tex[id];
        float o[4];
        oracle::texunit_sample(t.rgba, t.width, t.height, x, y, o);
        return make_float4(o[0], o[1], o[2], o[3]);
    }
};

struct HostScene {
    HostBvh bvh;
    std::vector<NrmRecord> nrms;
    std::vector<UvRecord> uvs;
    std::vector<MaterialRecord> mats;
    std::vector<std::vector<uint8_t>> tex_pixels;
    std::vector<B2ptTexture> textures;
    DeviceScene view;
};

template <class Rng>
float3 trace_path(const HostScene& hs, const CameraParams& cam, int x, int y, uint32_t frame_id, uint64_t seed,
                  const B2ptPointLight* lights, int n_lights, int max_bounces, int mode, uint64_t* rays) {
    PathVertex pv;
    pv.beta = v3(1.0f);
    pv.radiance = v3(0.0f);
    camera_ray(cam, x, y, pv.next_o, pv.next_d);
    RngCtx ctx;
    ctx.seed = seed;
    ctx.pixel = (uint32_t)(cam.width * y + x);
    ctx.sample = frame_id;
    ctx.lcg_state = RngIO<Rng>::first_state(ctx.pixel, frame_id);
    const HostTexFetch tex{hs.textures.data()};
    const AlphaCutout<HostTexFetch> cutout{hs.view, tex};
    for (int bounce = 0; path_continues(pv.beta, bounce, max_bounces); bounce++) {
        ctx.bounce = (uint32_t)bounce;
        HitRecord hit;
        rays[0]++;
        if (!traverse<false, false>(hs.view.nodes, hs.view.tris, hs.view.n_nodes, pv.next_o, pv.next_d, 0.0f, 100.0f,
                                    hit, nullptr, cutout))
            break;   // __miss__radiance
        const float3 ray_d = pv.next_d;
        SurfacePoint sp;
        {
            Rng rng = RngIO<Rng>::open(ctx, 0);
            surface_stage(hs.view, tex, hit, ray_d, lights, n_lights, rng, sp);
            ctx.lcg_state = RngIO<Rng>::close(rng);
        }
        bool visible = false;
        if (sp.light >= 0) {
            HitRecord sh;
            rays[1]++;
            visible = !traverse<true, false>(hs.view.nodes, hs.view.tris, hs.view.n_nodes, sp.sh_o, sp.sh_d, 0.0f,
                                             sp.sh_tmax, sh, nullptr, cutout);
        }
        int kind = mode - 1;
        if (mode == B2PT_BSDF_AUTO) kind = sp.conductor ? kBsdfConductor : kBsdfLayered;
        Rng rng = RngIO<Rng>::open(ctx, 1);
        bool alive = false;
        switch (kind) {
            case kBsdfConductor: alive = shade_stage<kBsdfConductor>(sp, visible, ray_d, lights, n_lights, rng, pv); break;
            case kBsdfDiffuse: alive = shade_stage<kBsdfDiffuse>(sp, visible, ray_d, lights, n_lights, rng, pv); break;
            case kBsdfDielectric: alive = shade_stage<kBsdfDielectric>(sp, visible, ray_d, lights, n_lights, rng, pv); break;
            default: alive = shade_stage<kBsdfLayered>(sp, visible, ray_d, lights, n_lights, rng, pv); break;
        }
        ctx.lcg_state = RngIO<Rng>::close(rng);
        if (!alive) break;
    }
    return pv.radiance;
}

}  // namespace

extern "C" {

void* hostsim_scene_create(const B2ptSceneDesc* desc) {
    HostScene* hs = new HostScene();
    std::vector<float> soup;
    std::vector<uint32_t> tri_mesh;
    std::vector<int32_t> tri_local;
    bool any_uv = false;
    for (int m = 0; m < desc->n_meshes; m++) {
        const B2ptMesh& me = desc->meshes[m];
        any_uv |= me.texcoords != nullptr;
        for (int t = 0; t < me.n_triangles; t++) {
            for (int k = 0; k < 3; k++) {
                const float* p = me.positions + 3 * me.indices[3 * t + k];
                const float4 w = mat4_mul(me.model, p[0], p[1], p[2], 1.0f);   // k_flatten
                soup.push_back(w.x); soup.push_back(w.y); soup.push_back(w.z);
            }
            tri_mesh.push_back((uint32_t)m);
            tri_local.push_back(t);
        }
        MaterialRecord mr;
        memcpy(mr.albedo, me.albedo, 12);
        mr.metallic = me.metallic;
        mr.roughness = me.roughness;
        auto ok = [&](int t) { return (t >= 0 && t < desc->n_textures) ? t : -1; };
        mr.tex_albedo = ok(me.tex_albedo);
        mr.tex_normal = ok(me.tex_normal);
        mr.tex_metal_rough = ok(me.tex_metal_rough);
        memcpy(mr.mcol, me.model, sizeof mr.mcol);
        hs->mats.push_back(mr);
    }
    const int n = (int)tri_mesh.size();
    build(soup.data(), n, hs->bvh);
    hs->nrms.resize(n);
    if (any_uv) hs->uvs.resize(n);
    for (int s = 0; s < n; s++) {
        const uint32_t g = hs->bvh.order[s];
        const B2ptMesh& me = desc->meshes[tri_mesh[g]];
        hs->bvh.tris[s].c.z = u2f_bits(tri_mesh[g]);   // mesh id next to the primitive id
        const int32_t* ix = me.indices + 3 * tri_local[g];
        float nn[9] = {0};
        if (me.normals)
            for (int k = 0; k < 3; k++) memcpy(nn + 3 * k, me.normals + 3 * ix[k], 12);
        NrmRecord nr;
        nr.a = make_float4(nn[0], nn[1], nn[2], nn[3]);
        nr.b = make_float4(nn[4], nn[5], nn[6], nn[7]);
        nr.c = make_float4(nn[8], u2f_bits(me.normals ? 1u : 0u), 0.0f, 0.0f);
        hs->nrms[s] = nr;
        if (any_uv) {
            UvRecord ur;
            ur.t0 = ur.t1 = ur.t2 = make_float2(0.0f, 0.0f);
            if (me.texcoords) {
                ur.t0 = make_float2(me.texcoords[2 * ix[0]], me.texcoords[2 * ix[0] + 1]);
                ur.t1 = make_float2(me.texcoords[2 * ix[1]], me.texcoords[2 * ix[1] + 1]);
                ur.t2 = make_float2(me.texcoords[2 * ix[2]], me.texcoords[2 * ix[2] + 1]);
            }
            hs->uvs[s] = ur;
        }
    }
    for (int i = 0; i < desc->n_textures; i++) {
        const B2ptTexture& t = desc->textures[i];
        hs->tex_pixels.emplace_back(t.rgba, t.rgba + (size_t)t.width * t.height * 4);
        hs->textures.push_back(B2ptTexture{hs->tex_pixels.back().data(), t.width, t.height});
    }
    DeviceScene& v = hs->view;
    memset(&v, 0, sizeof v);
    v.nodes = hs->bvh.nodes.data();
    v.tris = hs->bvh.tris.data();
    v.nrms = hs->nrms.data();
    v.uvs = any_uv ? hs->uvs.data() : nullptr;
    v.materials = hs->mats.data();
    v.n_nodes = (int)hs->bvh.nodes.size();
    v.n_tris = n;
    v.n_meshes = desc->n_meshes;
    v.n_textures = desc->n_textures;
    return hs;
}

void hostsim_scene_destroy(void* h) { delete static_cast<HostScene*>(h); }

// frames_out: n_frames * H * W * 3; rays: [radiance, shadow]
void hostsim_render(void* h, int width, int height, const float* cam_pos, const float* inv_view,
                    const float* inv_proj, const B2ptPointLight* lights, int n_lights, int max_bounces, int mode,
                    int rng_mode, uint64_t seed, uint32_t frame_first, int n_frames, float* frames_out,
                    uint64_t* rays_out) {
    const HostScene& hs = *static_cast<HostScene*>(h);
    CameraParams cam;
    cam.position = v3(cam_pos[0], cam_pos[1], cam_pos[2]);
    memcpy(cam.inv_view, inv_view, 64);
    memcpy(cam.inv_proj, inv_proj, 64);
    cam.width = width;
    cam.height = height;
    uint64_t r0 = 0, r1 = 0;
    for (int f = 0; f < n_frames; f++) {
        const uint32_t fid = frame_first + (uint32_t)f;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : r0, r1)
        for (int y = 0; y < height; y++)
            for (int x = 0; x < width; x++) {
                uint64_t rays[2] = {0, 0};
                float3 c;
                if (rng_mode == B2PT_RNG_LCG_HOST)
                    c = trace_path<LcgRng<true>>(hs, cam, x, y, fid, seed, lights, n_lights, max_bounces, mode, rays);
                else if (rng_mode == B2PT_RNG_LCG_DEVICE)
                    c = trace_path<LcgRng<false>>(hs, cam, x, y, fid, seed, lights, n_lights, max_bounces, mode, rays);
                else
                    c = trace_path<PhiloxRng>(hs, cam, x, y, fid, seed, lights, n_lights, max_bounces, mode, rays);
                float* o = frames_out + ((size_t)f * width * height + (size_t)y * width + x) * 3;
                o[0] = c.x; o[1] = c.y; o[2] = c.z;
                r0 += rays[0];
                r1 += rays[1];
            }
    }
    if (rays_out) { rays_out[0] = r0; rays_out[1] = r1; }
}

}  // extern "C"
