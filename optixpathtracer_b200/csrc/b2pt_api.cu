// b2pt_api.cu — the extern "C" surface declared in include/b2pt.h.
//
// Each function cites the member of the reference's OptixRenderer it replaces
// (OptixPathtracer/source/Renderer/OptiX/OptixRenderer.{h,cpp}).  Library code
// throws b2pt::Error internally; every entry point catches at the boundary and
// turns it into a status code + message.
#include <cmath>
#include <cstring>
#include <new>

#include "renderer.h"
#include "render.h"

using namespace b2pt;

namespace {

thread_local std::string g_create_error;

template <typename F>
int guarded(B2ptRenderer* r, F&& f) {
    try {
        if (r) B2PT_CUDA(cudaSetDevice(r->device));
        f();
        return B2PT_OK;
    } catch (const Error& e) {
        if (r) r->set_error(e.msg); else g_create_error = e.msg;
        return e.code;
    } catch (const std::bad_alloc&) {
        if (r) r->set_error("host allocation failed"); else g_create_error = "host allocation failed";
        return B2PT_E_NOMEM;
    } catch (const std::exception& e) {
        if (r) r->set_error(e.what()); else g_create_error = e.what();
        return B2PT_E_INVALID;
    }
}

LaunchCfg launch_cfg(B2ptRenderer* r) {
    LaunchCfg lc;
    lc.stream = r->stream;
    lc.grid = r->sm_count * trace_grid_blocks_per_sm();
    lc.stats = r->cfg.collect_stats != 0;
    lc.launch_counter = &r->kernel_launches;
    return lc;
}

// split n*3 interleaved floats into three planes
void deinterleave3(const float* in, int64_t n, std::vector<float>& out) {
    out.resize((size_t)n * 3);
    for (int64_t i = 0; i < n; i++) {
        out[i] = in[3 * i];
        out[n + i] = in[3 * i + 1];
        out[2 * n + i] = in[3 * i + 2];
    }
}

}  // namespace

extern "C" {

int b2pt_abi_version(void) { return B2PT_ABI_VERSION; }

void b2pt_default_config(B2ptConfig* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof *cfg);
    cfg->device = 0;                       // OptixRenderer.cpp:70
    cfg->bsdf_mode = B2PT_BSDF_AUTO;       // devicePrograms.cu:310-318
    cfg->rng_mode = B2PT_RNG_PHILOX;
    cfg->seed = 0x5EEDull;
    cfg->max_bounces = 8;
    cfg->wave_paths_log2 = 0;
    cfg->bvh_leaf_size = 0;
    cfg->collect_stats = 0;
}

int b2pt_create(const B2ptSceneDesc* scene, const B2ptConfig* cfg, B2ptRenderer** out) {
    if (!scene || !out) { g_create_error = "null argument"; return B2PT_E_INVALID; }
    *out = nullptr;
    B2ptRenderer* r = new (std::nothrow) B2ptRenderer();
    if (!r) { g_create_error = "host allocation failed"; return B2PT_E_NOMEM; }
    if (cfg) r->cfg = *cfg; else b2pt_default_config(&r->cfg);
    r->device = r->cfg.device;
    int rc = guarded(nullptr, [&] {
        // the same ranges the setters enforce: the per-bounce counters are laid out for at most 99 bounces and
        // an unknown BSDF mode would match no shade queue (its paths would never write their result)
        const B2ptConfig& c = r->cfg;
        B2PT_REQUIRE(c.max_bounces >= 0 && c.max_bounces <= 99, B2PT_E_INVALID, "max_bounces must be in [0, 99]");
        B2PT_REQUIRE(c.bsdf_mode >= B2PT_BSDF_AUTO && c.bsdf_mode <= B2PT_BSDF_LAYERED, B2PT_E_INVALID, "bad bsdf mode");
        B2PT_REQUIRE(c.rng_mode >= B2PT_RNG_PHILOX && c.rng_mode <= B2PT_RNG_LCG_HOST, B2PT_E_INVALID, "bad rng mode");
        B2PT_REQUIRE(c.wave_paths_log2 == 0 || (c.wave_paths_log2 >= 10 && c.wave_paths_log2 <= 27), B2PT_E_INVALID,
                     "wave_paths_log2 must be 0 (auto) or in [10, 27]");
        B2PT_REQUIRE(c.bvh_leaf_size >= 0 && c.bvh_leaf_size <= 8, B2PT_E_INVALID, "bvh_leaf_size must be in [0, 8]");
        int n_dev = 0;
        B2PT_CUDA(cudaGetDeviceCount(&n_dev));
        B2PT_REQUIRE(r->device >= 0 && r->device < n_dev, B2PT_E_INVALID, "device ordinal out of range");
        B2PT_CUDA(cudaSetDevice(r->device));
        B2PT_CUDA(cudaDeviceGetAttribute(&r->sm_count, cudaDevAttrMultiProcessorCount, r->device));
        B2PT_CUDA(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
        B2PT_CUDA(cudaEventCreate(&r->ev0));
        B2PT_CUDA(cudaEventCreate(&r->ev1));
        build_scene(*scene, r->scene, r->stream, r->cfg.bvh_leaf_size);
        r->d_counters.alloc(1);
        r->d_counters.zero(r->stream);
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
    if (rc != B2PT_OK) {
        b2pt_destroy(r);
        return rc;
    }
    *out = r;
    return B2PT_OK;
}

void b2pt_destroy(B2ptRenderer* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    if (r->dist_comm) b2pt_dist_shutdown(r);
    r->scene.free_textures();
    for (cudaEvent_t e : r->timing_events) cudaEventDestroy(e);
    if (r->ev0) cudaEventDestroy(r->ev0);
    if (r->ev1) cudaEventDestroy(r->ev1);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;  // DeviceBuffer members free their memory
}

const char* b2pt_last_error(const B2ptRenderer* r) {
    return r ? r->last_error.c_str() : g_create_error.c_str();
}

int b2pt_resize(B2ptRenderer* r, int32_t width, int32_t height) {
    if (!r) return B2PT_E_INVALID;
    if (width == 0 || height == 0) return B2PT_OK;  // minimised window: OptixRenderer.cpp:651
    if (width < 0 || height < 0) { r->set_error("negative size"); return B2PT_E_INVALID; }
    return guarded(r, [&] {
        B2PT_REQUIRE((int64_t)width * height < (1ll << 28), B2PT_E_UNSUPPORTED, "frame too large");
        r->width = width;
        r->height = height;
        r->d_accum.alloc((size_t)width * height * 3);
        r->d_accum.zero(r->stream);
        r->d_frame.alloc((size_t)width * height * 3);
        r->spp_accumulated = 0;
        r->wave_capacity = 0;  // work buffers are re-laid out on the next render
    });
}

int b2pt_set_camera(B2ptRenderer* r, const float position[3], const float inverse_view[16],
                    const float inverse_projection[16]) {
    if (!r || !position || !inverse_view || !inverse_projection) return B2PT_E_INVALID;
    memcpy(r->cam_pos, position, 12);
    memcpy(r->cam_inv_view, inverse_view, 64);
    memcpy(r->cam_inv_proj, inverse_projection, 64);
    r->have_camera = true;
    return B2PT_OK;
}

int b2pt_set_lights(B2ptRenderer* r, const B2ptPointLight* lights, int32_t n_lights) {
    if (!r || n_lights < 0 || (n_lights > 0 && !lights)) return B2PT_E_INVALID;
    if (n_lights > 255) { r->set_error("at most 255 point lights"); return B2PT_E_UNSUPPORTED; }
    return guarded(r, [&] {
        r->lights.assign(lights, lights + n_lights);
        r->d_lights.upload(r->lights.data(), r->lights.size(), r->stream);
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

int b2pt_set_max_bounces(B2ptRenderer* r, int32_t max_bounces) {
    if (!r) return B2PT_E_INVALID;
    if (max_bounces < 0 || max_bounces > 99) {  // __miss__radiance parks paths at bounce 100 (:582)
        r->set_error("max_bounces must be in [0, 99]");
        return B2PT_E_INVALID;
    }
    r->cfg.max_bounces = max_bounces;
    return B2PT_OK;
}

int b2pt_set_bsdf_mode(B2ptRenderer* r, int32_t mode) {
    if (!r) return B2PT_E_INVALID;
    if (mode < B2PT_BSDF_AUTO || mode > B2PT_BSDF_LAYERED) { r->set_error("bad bsdf mode"); return B2PT_E_INVALID; }
    r->cfg.bsdf_mode = mode;
    return B2PT_OK;
}

int b2pt_set_rng(B2ptRenderer* r, int32_t mode, uint64_t seed) {
    if (!r) return B2PT_E_INVALID;
    if (mode < B2PT_RNG_PHILOX || mode > B2PT_RNG_LCG_HOST) { r->set_error("bad rng mode"); return B2PT_E_INVALID; }
    r->cfg.rng_mode = mode;
    r->cfg.seed = seed;
    return B2PT_OK;
}

int b2pt_set_frame_id(B2ptRenderer* r, uint32_t frame_id) {
    if (!r) return B2PT_E_INVALID;
    r->frame_id = frame_id;
    return B2PT_OK;
}

int b2pt_get_frame_id(const B2ptRenderer* r, uint32_t* frame_id) {
    if (!r || !frame_id) return B2PT_E_INVALID;
    *frame_id = r->frame_id;
    return B2PT_OK;
}

// ---- rendering -------------------------------------------------------------

int b2pt_render(B2ptRenderer* r, float* h_rgb) {
    if (!r) return B2PT_E_INVALID;
    if (r->width == 0) return B2PT_OK;  // before the first resize: OptixRenderer.cpp:621
    if (!h_rgb) { r->set_error("null output buffer"); return B2PT_E_INVALID; }
    return guarded(r, [&] {
        render_frames(r, 1, /*accumulate=*/false);
        r->d_frame.download(h_rgb, (size_t)r->width * r->height * 3, r->stream);
        B2PT_CUDA(cudaStreamSynchronize(r->stream));  // CUDA_SYNC_CHECK + download, :642-644
    });
}

int b2pt_render_accumulate_device(B2ptRenderer* r, int32_t spp) {
    if (!r) return B2PT_E_INVALID;
    if (r->width == 0) return B2PT_OK;
    if (spp < 0) { r->set_error("negative spp"); return B2PT_E_INVALID; }
    return guarded(r, [&] {
        if (spp > 0) render_frames(r, spp, /*accumulate=*/true);
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

int b2pt_render_accumulate(B2ptRenderer* r, int32_t spp, float* h_rgb_mean) {
    if (!r) return B2PT_E_INVALID;
    if (r->width == 0) return B2PT_OK;
    if (spp < 0) { r->set_error("negative spp"); return B2PT_E_INVALID; }
    return guarded(r, [&] {
        if (spp > 0) render_frames(r, spp, /*accumulate=*/true);
        if (h_rgb_mean) {
            resolve_mean(r);  // d_frame = d_accum / spp_accumulated
            r->d_frame.download(h_rgb_mean, (size_t)r->width * r->height * 3, r->stream);
        }
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

int b2pt_clear_accumulation(B2ptRenderer* r) {
    if (!r) return B2PT_E_INVALID;
    return guarded(r, [&] {
        r->d_accum.zero(r->stream);
        r->spp_accumulated = 0;
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

int b2pt_get_accumulation_device(B2ptRenderer* r, float** d_rgb_sum, uint32_t* spp_accumulated) {
    if (!r) return B2PT_E_INVALID;
    if (d_rgb_sum) *d_rgb_sum = r->d_accum.get();
    if (spp_accumulated) *spp_accumulated = r->spp_accumulated;
    return B2PT_OK;
}

int b2pt_download_accumulation(B2ptRenderer* r, float* h_rgb_sum, uint32_t* spp_accumulated) {
    if (!r) return B2PT_E_INVALID;
    return guarded(r, [&] {
        if (h_rgb_sum) r->d_accum.download(h_rgb_sum, r->d_accum.size(), r->stream);
        if (spp_accumulated) *spp_accumulated = r->spp_accumulated;
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

int b2pt_upload_accumulation(B2ptRenderer* r, const float* h_rgb_sum, uint32_t spp_accumulated) {
    if (!r || !h_rgb_sum) return B2PT_E_INVALID;
    return guarded(r, [&] {
        B2PT_REQUIRE(r->width > 0, B2PT_E_STATE, "resize first");
        B2PT_CUDA(cudaMemcpyAsync(r->d_accum.get(), h_rgb_sum, r->d_accum.bytes(), cudaMemcpyHostToDevice, r->stream));
        r->spp_accumulated = spp_accumulated;
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

// ---- image output ----------------------------------------------------------

int b2pt_display_rgb8(B2ptRenderer* r, int32_t apply_oetf, int32_t top_row_first, uint8_t* h_rgb8) {
    if (!r || !h_rgb8) return B2PT_E_INVALID;
    return guarded(r, [&] { display_to_host(r, apply_oetf != 0, top_row_first != 0, h_rgb8); });
}
int b2pt_write_bmp(B2ptRenderer* r, const char* path, int32_t apply_oetf) {
    if (!r || !path) return B2PT_E_INVALID;
    return guarded(r, [&] { write_bmp(r, path, apply_oetf != 0); });
}
int b2pt_write_png(B2ptRenderer* r, const char* path, int32_t apply_oetf) {
    if (!r || !path) return B2PT_E_INVALID;
    return guarded(r, [&] { write_png(r, path, apply_oetf != 0); });
}
int b2pt_write_exr(B2ptRenderer* r, const char* path) {
    if (!r || !path) return B2PT_E_INVALID;
    return guarded(r, [&] { write_exr(r, path); });
}

// ---- statistics ------------------------------------------------------------

int b2pt_get_stats(const B2ptRenderer* rc, B2ptStats* out) {
    if (!rc || !out) return B2PT_E_INVALID;
    B2ptRenderer* r = const_cast<B2ptRenderer*>(rc);
    return guarded(r, [&] {
        TraceCounters c;
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
        B2PT_CUDA(cudaMemcpy(&c, r->d_counters.get(), sizeof c, cudaMemcpyDeviceToHost));
        memset(out, 0, sizeof *out);
        out->samples = r->samples;
        out->radiance_rays = c.radiance_rays + r->hook_radiance_rays;
        out->shadow_rays = c.shadow_rays + r->hook_shadow_rays;
        out->nodes_visited = c.nodes;
        out->triangles_tested = c.tris;
        out->exact_tests = c.exact;
        out->nan_pixels = c.nan_samples;
        out->kernel_launches = r->kernel_launches;
        out->device_ms_render = r->device_ms_render;
        out->device_ms_build = r->scene.build_ms;
        out->n_triangles = r->scene.n_tris;
        out->n_nodes = r->scene.n_nodes;
        out->node_bytes = trace_node_bytes(r->scene.view());
        out->triangle_bytes = (int32_t)sizeof(TriRecord);
        out->scene_device_bytes = r->scene.device_bytes();
        out->nodes_visited_shadow = c.nodes_any;
        out->triangles_tested_shadow = c.tris_any;
        out->exact_tests_shadow = c.exact_any;
        for (int k = 0; k < 5; k++) {
            out->device_ms_kernel[k] = r->ms_kernel[k];
            out->launches_kernel[k] = r->launches_kernel_cls[k];
        }
        out->device_ms_kernel[kClsShade] += r->ms_kernel[kClsEngine];   // the engine is a shade kernel
        out->launches_kernel[kClsShade] += r->launches_kernel_cls[kClsEngine];
        out->device_ms_engine = r->ms_kernel[kClsEngine];
        out->launches_engine = r->launches_kernel_cls[kClsEngine];
        out->n_wide_nodes = r->scene.n_wide_nodes;
        out->bvh_depth = r->scene.depth;
        out->bvh_wide_depth = r->scene.wide_depth;
        out->walk_f_hits = c.walk_f_hits;
        out->walk_s_hits = c.walk_s_hits;
        out->shadow_rays_traced = out->shadow_rays - c.shadow_culled;
        out->warp_slots_node = c.node_slots;
        out->warp_slots_tri = c.tri_slots;
    });
}

int b2pt_set_kernel_timing(B2ptRenderer* r, int32_t on) {
    if (!r) return B2PT_E_INVALID;
    r->kernel_timing = on != 0;
    return B2PT_OK;
}

int b2pt_reset_stats(B2ptRenderer* r) {
    if (!r) return B2PT_E_INVALID;
    return guarded(r, [&] {
        r->d_counters.zero(r->stream);
        r->samples = 0;
        r->kernel_launches = 0;
        r->hook_radiance_rays = r->hook_shadow_rays = 0;
        r->device_ms_render = 0.0;
        for (int k = 0; k < kClsCount; k++) { r->ms_kernel[k] = 0.0; r->launches_kernel_cls[k] = 0; }
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

// ---- parity hooks: the renderer's own K5/K8 kernels on caller rays ----------

int b2pt_trace_closest(B2ptRenderer* r, int64_t n, const float* org, const float* dir, float tmin,
                       float tmax, int32_t* prim, int32_t* mesh, float* t, float* u, float* v) {
    if (!r || n < 0 || (n > 0 && (!org || !dir))) return B2PT_E_INVALID;
    if (n == 0) return B2PT_OK;
    return guarded(r, [&] {
        B2PT_REQUIRE(n < (1ll << 31), B2PT_E_UNSUPPORTED, "too many rays in one call");
        std::vector<float> ho, hd;
        deinterleave3(org, n, ho);
        deinterleave3(dir, n, hd);
        DeviceBuffer<float> d_o, d_d;
        DeviceBuffer<HitRecord> d_h;
        DeviceBuffer<uint32_t> d_n;
        d_o.upload(ho.data(), ho.size(), r->stream);
        d_d.upload(hd.data(), hd.size(), r->stream);
        d_h.alloc((size_t)n);
        const uint32_t n32 = (uint32_t)n;
        d_n.upload(&n32, 1, r->stream);
        PathState st = {};
        st.ox = d_o.get(); st.oy = d_o.get() + n; st.oz = d_o.get() + 2 * n;
        st.dx = d_d.get(); st.dy = d_d.get() + n; st.dz = d_d.get() + 2 * n;
        launch_closest(launch_cfg(r), r->scene.view(), st, d_n.get(), tmin, tmax, d_h.get(), r->d_counters.get());
        r->hook_radiance_rays += (uint64_t)n;
        B2PT_CUDA(cudaGetLastError());
        std::vector<HitRecord> hh((size_t)n);
        d_h.download(hh.data(), hh.size(), r->stream);
        std::vector<uint32_t> order((size_t)r->scene.n_tris);
        r->scene.tri_order.download(order.data(), order.size(), r->stream);
        std::vector<TriRecord> tris;
        if (mesh) {
            tris.resize((size_t)r->scene.n_tris);
            r->scene.tris.download(tris.data(), tris.size(), r->stream);
        }
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
        for (int64_t i = 0; i < n; i++) {
            const HitRecord& h = hh[i];
            const bool hit = h.tri != kHitMiss;
            const uint32_t slot = h.tri & ~kHitBackface;
            if (prim) prim[i] = hit ? (int32_t)order[slot] : -1;
            if (mesh) mesh[i] = hit ? (int32_t)f2u_bits(tris[slot].c.z) : -1;
            if (t) t[i] = hit ? h.t : 0.0f;
            if (u) u[i] = hit ? h.u : 0.0f;
            if (v) v[i] = hit ? h.v : 0.0f;
        }
    });
}

int b2pt_trace_any(B2ptRenderer* r, int64_t n, const float* org, const float* dir, float tmin,
                   const float* tmax, uint8_t* occluded) {
    if (!r || n < 0 || (n > 0 && (!org || !dir || !tmax || !occluded))) return B2PT_E_INVALID;
    if (n == 0) return B2PT_OK;
    return guarded(r, [&] {
        B2PT_REQUIRE(n < (1ll << 31), B2PT_E_UNSUPPORTED, "too many rays in one call");
        std::vector<float> ho, hd;
        deinterleave3(org, n, ho);
        deinterleave3(dir, n, hd);
        DeviceBuffer<float> d_o, d_d, d_t;
        DeviceBuffer<uint8_t> d_occ;
        DeviceBuffer<uint32_t> d_n;
        d_o.upload(ho.data(), ho.size(), r->stream);
        d_d.upload(hd.data(), hd.size(), r->stream);
        d_t.upload(tmax, (size_t)n, r->stream);
        d_occ.alloc((size_t)n);
        d_occ.zero(r->stream);
        const uint32_t n32 = (uint32_t)n;
        d_n.upload(&n32, 1, r->stream);
        ShadowRays sh;
        sh.ox = d_o.get(); sh.oy = d_o.get() + n; sh.oz = d_o.get() + 2 * n;
        sh.dx = d_d.get(); sh.dy = d_d.get() + n; sh.dz = d_d.get() + 2 * n;
        sh.tmax = d_t.get();
        launch_any(launch_cfg(r), r->scene.view(), sh, d_n.get(), tmin, nullptr, d_occ.get(), r->d_counters.get());
        r->hook_shadow_rays += (uint64_t)n;
        B2PT_CUDA(cudaGetLastError());
        d_occ.download(occluded, (size_t)n, r->stream);
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

int b2pt_download_bvh(B2ptRenderer* r, void* nodes, int32_t* tri_order) {
    if (!r) return B2PT_E_INVALID;
    return guarded(r, [&] {
        if (nodes) r->scene.nodes.download(static_cast<BvhNode2*>(nodes), r->scene.nodes.size(), r->stream);
        if (tri_order)
            r->scene.tri_order.download(reinterpret_cast<uint32_t*>(tri_order), r->scene.tri_order.size(), r->stream);
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

}  // extern "C"
