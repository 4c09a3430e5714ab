// noalloc.cu — b2pt_trace_any_noalloc: the occlusion hook (K8) for callers that cannot allocate.
//
// No counterpart in the reference.  b2pt_trace_any (b2pt_api.cu)
// cudaMallocs its ray planes, the ray count and the answer on every call, which code under CUDA-graph capture, a
// tight serving loop or a harness that forbids allocation cannot do.  This variant runs the SAME traversal
// kernel (launch_any, trace_kernels.cu) with all of that in ONE file-scope array of device memory:
//
//     [ header 256 B: ray count ][ org n*3 | dir n*3 as given ][ ox oy oz dx dy dz tmax planes ][ occluded ]
//
// sized for kNoAllocMaxRays rays (B2PT_NOALLOC_MAX_RAYS, include/b2pt.h): 1.7 MB that every GPU that loads the
// library holds until the process exits.  The entry point calls no cudaMalloc / cudaFree and no allocating
// library routine; larger calls are refused (B2PT_E_UNSUPPORTED).
//
// The array is shared by every call on a device and keeps the last call's contents: what must start at zero
// (the answers: rays with tmax < 0 are not traced and report 0) is cleared per call, everything else is
// overwritten before it is read.  Calls that overlap on one GPU (two renderers, two host threads) RACE on it;
// callers serialise them.
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "renderer.h"
#include "render.h"

namespace b2pt {

namespace {

constexpr int64_t kNoAllocMaxRays = B2PT_NOALLOC_MAX_RAYS;
constexpr size_t kHeaderBytes = 256;
constexpr size_t kAosBytes = (size_t)kNoAllocMaxRays * 6 * sizeof(float);      // org, dir as the caller holds them
constexpr size_t kPlaneBytes = (size_t)kNoAllocMaxRays * 7 * sizeof(float);    // ox oy oz dx dy dz tmax
constexpr size_t kOccBytes = (size_t)kNoAllocMaxRays;
constexpr size_t kScratchBytes = kHeaderBytes + kAosBytes + kPlaneBytes + kOccBytes;

// In host code this name is a host-side stand-in, NOT a device address: only scratch_of() below turns it into one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_lock;
unsigned char* g_scratch_ptr[kMaxDevices] = {};

// device address of the array on `device` (which must be current); resolved once per device
unsigned char* scratch_of(int device) {
    B2PT_REQUIRE(device >= 0 && device < kMaxDevices, B2PT_E_UNSUPPORTED, "b2pt_trace_any_noalloc: device index out of range");
    std::lock_guard<std::mutex> g(g_lock);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        B2PT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        B2PT_REQUIRE(p != nullptr, B2PT_E_CUDA, "b2pt_trace_any_noalloc: cudaGetSymbolAddress returned null");
        g_scratch_ptr[device] = static_cast<unsigned char*>(p);
    }
    return g_scratch_ptr[device];
}

// n*3 interleaved -> three planes of n, for org and dir; also publishes the ray count
__global__ void __launch_bounds__(256)
k_noalloc_planes(uint32_t n, const float* __restrict__ org, const float* __restrict__ dir, float* __restrict__ planes,
                 uint32_t* __restrict__ count) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *count = n;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        planes[i] = org[3 * i];
        planes[(size_t)n + i] = org[3 * i + 1];
        planes[2 * (size_t)n + i] = org[3 * i + 2];
        planes[3 * (size_t)n + i] = dir[3 * i];
        planes[4 * (size_t)n + i] = dir[3 * i + 1];
        planes[5 * (size_t)n + i] = dir[3 * i + 2];
    }
}

}  // namespace

}  // namespace b2pt

using namespace b2pt;

extern "C" int b2pt_trace_any_noalloc(B2ptRenderer* r, int64_t n, const float* org, const float* dir, float tmin,
                                      const float* tmax, uint8_t* occluded) {
    if (!r || n < 0 || (n > 0 && (!org || !dir || !tmax || !occluded))) return B2PT_E_INVALID;
    if (n == 0) return B2PT_OK;
    try {
        if (n > kNoAllocMaxRays) {
            char msg[160];
            snprintf(msg, sizeof msg, "b2pt_trace_any_noalloc: %lld rays exceed the static scratch (B2PT_NOALLOC_MAX_RAYS = %lld); "
                     "split the call or use b2pt_trace_any", (long long)n, (long long)kNoAllocMaxRays);
            throw Error{B2PT_E_UNSUPPORTED, msg};
        }
        B2PT_CUDA(cudaSetDevice(r->device));
        unsigned char* base = scratch_of(r->device);
        uint32_t* d_count = reinterpret_cast<uint32_t*>(base);
        float* d_org = reinterpret_cast<float*>(base + kHeaderBytes);
        float* d_dir = d_org + 3 * n;
        float* d_planes = reinterpret_cast<float*>(base + kHeaderBytes + kAosBytes);
        uint8_t* d_occ = base + kHeaderBytes + kAosBytes + kPlaneBytes;
        cudaStream_t s = r->stream;
        B2PT_CUDA(cudaMemcpyAsync(d_org, org, (size_t)n * 12, cudaMemcpyHostToDevice, s));
        B2PT_CUDA(cudaMemcpyAsync(d_dir, dir, (size_t)n * 12, cudaMemcpyHostToDevice, s));
        B2PT_CUDA(cudaMemcpyAsync(d_planes + 6 * n, tmax, (size_t)n * 4, cudaMemcpyHostToDevice, s));
        B2PT_CUDA(cudaMemsetAsync(d_occ, 0, (size_t)n, s));   // the last call's answers are still there
        const uint32_t n32 = (uint32_t)n;
        k_noalloc_planes<<<(unsigned)std::min<int64_t>((n + 255) / 256, 1024), 256, 0, s>>>(n32, d_org, d_dir, d_planes, d_count);
        ++r->kernel_launches;
        ShadowRays sh;
        sh.ox = d_planes; sh.oy = d_planes + n; sh.oz = d_planes + 2 * n;
        sh.dx = d_planes + 3 * n; sh.dy = d_planes + 4 * n; sh.dz = d_planes + 5 * n;
        sh.tmax = d_planes + 6 * n;
        LaunchCfg lc;
        lc.stream = s;
        lc.grid = r->sm_count * trace_grid_blocks_per_sm();
        lc.stats = r->cfg.collect_stats != 0;
        lc.launch_counter = &r->kernel_launches;
        launch_any(lc, r->scene.view(), sh, d_count, tmin, nullptr, d_occ, r->d_counters.get());
        r->hook_shadow_rays += (uint64_t)n;
        B2PT_CUDA(cudaGetLastError());
        B2PT_CUDA(cudaMemcpyAsync(occluded, d_occ, (size_t)n, cudaMemcpyDeviceToHost, s));
        B2PT_CUDA(cudaStreamSynchronize(s));
        return B2PT_OK;
    } catch (const Error& e) {
        r->set_error(e.msg);
        return e.code;
    } catch (const std::exception& e) {
        r->set_error(e.what());
        return B2PT_E_INVALID;
    }
}
