// multi.cu — the multi-GPU entry points of include/b2pt.h.
//
// The reference renders on device 0 only (OptixRenderer.cpp:70).  Every (pixel, sample) path is independent given
// the read-only scene, so the path shards by SAMPLES: device g renders a disjoint range of absolute sample
// indices (the Philox key holds the absolute index, so the union is the 1-GPU image up to fp32 re-association),
// the scene + BVH are replicated (built concurrently, 1-5 ms), and the only exchange is one sum of the W*H*3 fp32
// accumulation buffers at the end of a call.  Two ways to drive it:
//
//  * one PROCESS, N devices (b2pt_create_multi ...): one host thread per device renders, then ONE kernel on the
//    root device reads every peer's accumulation buffer through NVLink peer memory, adds them in device order
//    (deterministic), applies the 1 / spp scale and writes the frame that is downloaded — reduce and epilogue
//    fused, no staging copy, no library.  Devices without peer access are staged with cudaMemcpyPeerAsync.
//  * one process PER device (torchrun; b2pt_dist_*): ncclReduce(sum, fp32, root 0) on the renderer's own stream,
//    straight out of the accumulation buffer K9 wrote, then the root's scale kernel.  NCCL is dlopen'ed
//    (libnccl.so.2) when b2pt_dist_init is first called, so libb2pt.so itself loads without it.
#include <dlfcn.h>
#include <nccl.h>   // types and enums only; no NCCL symbol is linked

#include <algorithm>
#include <thread>

#include "render.h"

using namespace b2pt;

// ---- kernels -------------------------------------------------------------------------------------------------
namespace {

constexpr int kMaxDevices = 16;
struct PeerPtrs {
    const float* p[kMaxDevices];
};

// out[i] = scale * sum_g src[g][i]; src[g] may live on a peer device (NVLink loads)
__global__ void __launch_bounds__(256)
k_reduce_peers(PeerPtrs src, int n_src, size_t n4, size_t n, float scale, float* __restrict__ out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 a = reinterpret_cast<const float4*>(src.p[0])[i];
        for (int g = 1; g < n_src; g++) {
            const float4 b = reinterpret_cast<const float4*>(src.p[g])[i];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        reinterpret_cast<float4*>(out)[i] = make_float4(a.x * scale, a.y * scale, a.z * scale, a.w * scale);
    }
    for (size_t i = 4 * n4 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride) {
        float a = src.p[0][i];
        for (int g = 1; g < n_src; g++) a += src.p[g][i];
        out[i] = a * scale;
    }
}

__global__ void __launch_bounds__(256)
k_scale_by_count(size_t n, float* __restrict__ buf, const unsigned long long* __restrict__ spp_total) {
    const unsigned long long s = *spp_total;
    const float scale = s ? 1.0f / (float)s : 0.0f;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) buf[i] *= scale;
}

thread_local std::string g_multi_error;

}  // namespace

// ---- single process, N devices ----------------------------------------------------------------------------------
struct B2ptMulti {
    std::vector<B2ptRenderer*> r;
    std::vector<int> dev;
    std::vector<char> peer;                 // root can load from device g's memory
    std::vector<DeviceBuffer<float>> stage; // on the root, for devices without peer access
    std::string last_error;
};

namespace {

template <typename F>
int multi_guard(B2ptMulti* m, F&& f) {
    try {
        f();
        return B2PT_OK;
    } catch (const Error& e) {
        if (m) m->last_error = e.msg; else g_multi_error = e.msg;
        return e.code;
    } catch (const std::exception& e) {
        if (m) m->last_error = e.what(); else g_multi_error = e.what();
        return B2PT_E_INVALID;
    }
}

// run f(g) for every device on its own host thread; first failure wins
template <typename F>
void for_each_device(B2ptMulti* m, F&& f) {
    const int n = (int)m->r.size();
    std::vector<int> rc(n, B2PT_OK);
    std::vector<std::string> msg(n);
    std::vector<std::thread> th;
    for (int g = 0; g < n; g++)
        th.emplace_back([&, g] {
            try {
                B2PT_CUDA(cudaSetDevice(m->dev[g]));
                f(g);
            } catch (const Error& e) {
                rc[g] = e.code;
                msg[g] = e.msg;
            } catch (const std::exception& e) {
                rc[g] = B2PT_E_INVALID;
                msg[g] = e.what();
            }
        });
    for (auto& t : th) t.join();
    for (int g = 0; g < n; g++)
        if (rc[g] != B2PT_OK) throw Error{rc[g], "device " + std::to_string(m->dev[g]) + ": " + msg[g]};
}

}  // namespace

extern "C" {

int b2pt_create_multi(const int32_t* devices, int32_t n_devices, const B2ptSceneDesc* scene, const B2ptConfig* cfg,
                      B2ptMulti** out) {
    if (!devices || !scene || !out || n_devices < 1 || n_devices > kMaxDevices) {
        g_multi_error = "bad argument (1..16 devices)";
        return B2PT_E_INVALID;
    }
    *out = nullptr;
    B2ptMulti* m = new (std::nothrow) B2ptMulti();
    if (!m) return B2PT_E_NOMEM;
    m->r.assign(n_devices, nullptr);
    m->dev.assign(devices, devices + n_devices);
    // replicated scene: every device uploads and builds its own BVH, concurrently
    std::vector<int> rc(n_devices, B2PT_OK);
    std::vector<std::string> msg(n_devices);
    std::vector<std::thread> th;
    for (int g = 0; g < n_devices; g++)
        th.emplace_back([&, g] {
            B2ptConfig c;
            if (cfg) c = *cfg; else b2pt_default_config(&c);
            c.device = devices[g];
            rc[g] = b2pt_create(scene, &c, &m->r[g]);
            if (rc[g] != B2PT_OK) msg[g] = b2pt_last_error(nullptr);   // thread-local message of the failed create
        });
    for (auto& t : th) t.join();
    for (int g = 0; g < n_devices; g++)
        if (rc[g] != B2PT_OK) {
            g_multi_error = "device " + std::to_string(devices[g]) + ": " + msg[g];
            const int code = rc[g];
            for (B2ptRenderer* r : m->r) b2pt_destroy(r);
            delete m;
            return code;
        }
    // peer access from the root to everybody else
    m->peer.assign(n_devices, 0);
    m->stage.resize(n_devices);
    const int root = m->dev[0];
    cudaSetDevice(root);
    for (int g = 0; g < n_devices; g++) {
        if (m->dev[g] == root) { m->peer[g] = 1; continue; }
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, root, m->dev[g]) == cudaSuccess && can) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[g], 0);
            if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) m->peer[g] = 1;
            cudaGetLastError();   // clear "already enabled"
        }
    }
    *out = m;
    return B2PT_OK;
}

void b2pt_destroy_multi(B2ptMulti* m) {
    if (!m) return;
    for (B2ptRenderer* r : m->r) b2pt_destroy(r);
    cudaSetDevice(m->dev[0]);
    m->stage.clear();
    delete m;
}

const char* b2pt_multi_last_error(const B2ptMulti* m) { return m ? m->last_error.c_str() : g_multi_error.c_str(); }
int32_t b2pt_multi_device_count(const B2ptMulti* m) { return m ? (int32_t)m->r.size() : 0; }
B2ptRenderer* b2pt_multi_renderer(B2ptMulti* m, int32_t i) {
    return (m && i >= 0 && i < (int32_t)m->r.size()) ? m->r[i] : nullptr;
}

#define B2PT_BROADCAST(call)                                                          \
    if (!m) return B2PT_E_INVALID;                                                    \
    for (B2ptRenderer* r : m->r) {                                                    \
        const int rc = (call);                                                        \
        if (rc != B2PT_OK) { m->last_error = b2pt_last_error(r); return rc; }         \
    }                                                                                 \
    return B2PT_OK;

int b2pt_multi_resize(B2ptMulti* m, int32_t w, int32_t h) { B2PT_BROADCAST(b2pt_resize(r, w, h)) }
int b2pt_multi_set_camera(B2ptMulti* m, const float position[3], const float inverse_view[16], const float inverse_projection[16]) {
    B2PT_BROADCAST(b2pt_set_camera(r, position, inverse_view, inverse_projection))
}
int b2pt_multi_set_lights(B2ptMulti* m, const B2ptPointLight* lights, int32_t n) { B2PT_BROADCAST(b2pt_set_lights(r, lights, n)) }
int b2pt_multi_set_max_bounces(B2ptMulti* m, int32_t b) { B2PT_BROADCAST(b2pt_set_max_bounces(r, b)) }
int b2pt_multi_set_bsdf_mode(B2ptMulti* m, int32_t mode) { B2PT_BROADCAST(b2pt_set_bsdf_mode(r, mode)) }
int b2pt_multi_set_rng(B2ptMulti* m, int32_t mode, uint64_t seed) { B2PT_BROADCAST(b2pt_set_rng(r, mode, seed)) }
int b2pt_multi_set_frame_id(B2ptMulti* m, uint32_t id) { B2PT_BROADCAST(b2pt_set_frame_id(r, id)) }
int b2pt_multi_clear_accumulation(B2ptMulti* m) { B2PT_BROADCAST(b2pt_clear_accumulation(r)) }
#undef B2PT_BROADCAST

// `spp_total` further samples per pixel, split over the devices: device g takes the g-th contiguous share of the
// absolute sample indices frame_id + 1 .. frame_id + spp_total.  h_rgb_mean (may be NULL) receives
// (sum over devices of the accumulated sums) / (samples accumulated since the last clear).
int b2pt_multi_render_accumulate(B2ptMulti* m, int32_t spp_total, float* h_rgb_mean) {
    if (!m) return B2PT_E_INVALID;
    if (spp_total < 0) { m->last_error = "negative spp"; return B2PT_E_INVALID; }
    B2ptRenderer* root = m->r[0];
    if (root->width == 0) return B2PT_OK;   // before the first resize: silent no-op like OptixRenderer.cpp:621
    return multi_guard(m, [&] {
        const int n = (int)m->r.size();
        const uint32_t base = root->frame_id;
        std::vector<int> share(n), first(n);
        for (int g = 0, at = 0; g < n; g++) {
            share[g] = spp_total / n + (g < spp_total % n ? 1 : 0);
            first[g] = at;
            at += share[g];
        }
        for_each_device(m, [&](int g) {
            B2ptRenderer* r = m->r[g];
            r->frame_id = base + (uint32_t)first[g];
            if (share[g] > 0) render_frames(r, share[g], /*accumulate=*/true);
            B2PT_CUDA(cudaStreamSynchronize(r->stream));
            r->frame_id = base + (uint32_t)spp_total;   // every device continues after the whole call's samples
        });
        if (!h_rgb_mean) return;
        // fused reduce + epilogue on the root, peers read in place over NVLink
        B2PT_CUDA(cudaSetDevice(m->dev[0]));
        const size_t nf = (size_t)root->width * root->height * 3;
        PeerPtrs src;
        uint64_t spp_sum = 0;
        for (int g = 0; g < n; g++) {
            spp_sum += m->r[g]->spp_accumulated;
            if (m->peer[g]) {
                src.p[g] = m->r[g]->d_accum.get();
            } else {
                m->stage[g].alloc(nf);
                B2PT_CUDA(cudaMemcpyPeerAsync(m->stage[g].get(), m->dev[0], m->r[g]->d_accum.get(), m->dev[g], nf * 4, root->stream));
                src.p[g] = m->stage[g].get();
            }
        }
        const float scale = spp_sum ? 1.0f / (float)spp_sum : 0.0f;
        k_reduce_peers<<<root->sm_count * 4, 256, 0, root->stream>>>(src, n, nf / 4, nf, scale, root->d_frame.get());
        root->kernel_launches++;
        B2PT_CUDA(cudaGetLastError());
        root->d_frame.download(h_rgb_mean, nf, root->stream);
        B2PT_CUDA(cudaStreamSynchronize(root->stream));
    });
}

}  // extern "C"

// ---- one process per device: NCCL ---------------------------------------------------------------------------------
namespace {

struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclReduce) Reduce = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
};

NcclApi& nccl() {
    static NcclApi api = [] {
        NcclApi a;
        // RTLD_NOLOAD first: a host program (e.g. PyTorch) may already hold the NCCL it wants used
        a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!a.lib) a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!a.lib) return a;
#define B2PT_SYM(name) a.name = reinterpret_cast<decltype(a.name)>(dlsym(a.lib, "nccl" #name))
        B2PT_SYM(GetUniqueId); B2PT_SYM(CommInitRank); B2PT_SYM(CommDestroy); B2PT_SYM(Reduce);
        B2PT_SYM(GroupStart); B2PT_SYM(GroupEnd); B2PT_SYM(GetErrorString); B2PT_SYM(GetVersion);
#undef B2PT_SYM
        return a;
    }();
    return api;
}

void require_nccl() {
    const NcclApi& a = nccl();
    B2PT_REQUIRE(a.lib && a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.Reduce && a.GroupStart && a.GroupEnd,
                 B2PT_E_UNSUPPORTED, "libnccl.so.2 could not be loaded (needed only by the b2pt_dist_* entry points)");
}

#define B2PT_NCCL(call)                                                                                    \
    do {                                                                                                   \
        ncclResult_t e_ = (call);                                                                          \
        if (e_ != ncclSuccess) {                                                                           \
            const char* s_ = nccl().GetErrorString ? nccl().GetErrorString(e_) : "?";                      \
            throw ::b2pt::Error{B2PT_E_CUDA, std::string(#call) + " failed: " + s_};                       \
        }                                                                                                  \
    } while (0)

template <typename F>
int dist_guard(B2ptRenderer* r, F&& f) {
    try {
        if (r) B2PT_CUDA(cudaSetDevice(r->device));
        f();
        return B2PT_OK;
    } catch (const Error& e) {
        if (r) r->set_error(e.msg); else g_multi_error = e.msg;
        return e.code;
    } catch (const std::exception& e) {
        if (r) r->set_error(e.what()); else g_multi_error = e.what();
        return B2PT_E_INVALID;
    }
}

}  // namespace

extern "C" {

int b2pt_dist_unique_id(uint8_t id[B2PT_DIST_ID_BYTES]) {
    if (!id) return B2PT_E_INVALID;
    return dist_guard(nullptr, [&] {
        require_nccl();
        static_assert(sizeof(ncclUniqueId) == B2PT_DIST_ID_BYTES, "ncclUniqueId is 128 bytes");
        ncclUniqueId u;
        B2PT_NCCL(nccl().GetUniqueId(&u));
        memcpy(id, &u, sizeof u);
    });
}

int b2pt_dist_init(B2ptRenderer* r, const uint8_t id[B2PT_DIST_ID_BYTES], int32_t rank, int32_t world) {
    if (!r || !id || world < 1 || rank < 0 || rank >= world) return B2PT_E_INVALID;
    return dist_guard(r, [&] {
        require_nccl();
        B2PT_REQUIRE(!r->dist_comm, B2PT_E_STATE, "b2pt_dist_init was already called on this renderer");
        ncclUniqueId u;
        memcpy(&u, id, sizeof u);
        ncclComm_t comm = nullptr;
        B2PT_NCCL(nccl().CommInitRank(&comm, world, u, rank));
        r->dist_comm = comm;
        r->dist_rank = rank;
        r->dist_world = world;
        r->d_dist_count.alloc(2);
    });
}

int b2pt_dist_shutdown(B2ptRenderer* r) {
    if (!r) return B2PT_E_INVALID;
    return dist_guard(r, [&] {
        if (r->dist_comm) {
            B2PT_CUDA(cudaStreamSynchronize(r->stream));
            B2PT_NCCL(nccl().CommDestroy(static_cast<ncclComm_t>(r->dist_comm)));
            r->dist_comm = nullptr;
        }
    });
}

// This rank traces `spp` further samples per pixel starting at ITS frame id (the caller hands every rank a disjoint
// range with b2pt_set_frame_id), the accumulation buffers of all ranks are summed on rank 0 with one ncclReduce on
// the renderer's stream, and rank 0's h_rgb_mean (may be NULL; ignored on the other ranks) receives
// sum / (samples accumulated over all ranks since their last clear).  Collective: every rank must call it.
int b2pt_dist_render_accumulate(B2ptRenderer* r, int32_t spp, float* h_rgb_mean) {
    if (!r) return B2PT_E_INVALID;
    if (spp < 0) { r->set_error("negative spp"); return B2PT_E_INVALID; }
    return dist_guard(r, [&] {
        B2PT_REQUIRE(r->dist_comm, B2PT_E_STATE, "b2pt_dist_init first");
        B2PT_REQUIRE(r->width > 0, B2PT_E_STATE, "resize first");
        if (spp > 0) render_frames(r, spp, /*accumulate=*/true);
        const size_t nf = (size_t)r->width * r->height * 3;
        const unsigned long long mine = r->spp_accumulated;
        unsigned long long* cnt = r->d_dist_count.get();
        B2PT_CUDA(cudaMemcpyAsync(cnt, &mine, 8, cudaMemcpyHostToDevice, r->stream));
        ncclComm_t comm = static_cast<ncclComm_t>(r->dist_comm);
        B2PT_NCCL(nccl().GroupStart());
        B2PT_NCCL(nccl().Reduce(r->d_accum.get(), r->d_frame.get(), nf, ncclFloat32, ncclSum, 0, comm, r->stream));
        B2PT_NCCL(nccl().Reduce(cnt, cnt + 1, 1, ncclUint64, ncclSum, 0, comm, r->stream));
        B2PT_NCCL(nccl().GroupEnd());
        if (r->dist_rank == 0) {
            k_scale_by_count<<<r->sm_count * 4, 256, 0, r->stream>>>(nf, r->d_frame.get(), cnt + 1);   // the root's epilogue
            r->kernel_launches++;
            B2PT_CUDA(cudaGetLastError());
            if (h_rgb_mean) r->d_frame.download(h_rgb_mean, nf, r->stream);
        }
        B2PT_CUDA(cudaStreamSynchronize(r->stream));
    });
}

}  // extern "C"
