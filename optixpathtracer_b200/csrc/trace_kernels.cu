// trace_kernels.cu — K5 (closest hit) and K8 (any hit) traversal kernels.
//
// K5 replaces TraceRadiance / optixTrace(RADIANCE) (devicePrograms.cu:243-260),
// K8 replaces LightVisibility / optixTrace(SHADOW, TERMINATE_ON_FIRST_HIT)
// (:216-241) with __miss__shadow's "visible = 1" (:585-591) folded in.
// Both are persistent grid-stride kernels: the number of live rays is read
// from device memory, so the host never synchronises inside the bounce loop.
#include "wavefront.cuh"
#include "traverse.cuh"

namespace b2pt {

namespace {

constexpr int kTraceThreads = 128;

template <bool STATS>
__global__ void __launch_bounds__(kTraceThreads)
k_closest(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, float tmin, float tmax,
          HitRecord* __restrict__ hits, TraceCounters* counters) {
    const uint32_t n = *n_active;
    TraceStats ts = {0, 0, 0};
    uint32_t rays = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float3 o = v3(st.ox[i], st.oy[i], st.oz[i]);
        const float3 d = v3(st.dx[i], st.dy[i], st.dz[i]);
        HitRecord h;
        traverse<false, STATS>(sc.nodes, sc.tris, sc.n_nodes, o, d, tmin, tmax, h, &ts);
        reinterpret_cast<float4*>(hits)[i] =
            make_float4(__uint_as_float(h.tri), h.t, h.u, h.v);
        rays++;
    }
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->radiance_rays, (unsigned long long)rays);
    }
}

template <bool STATS>
__global__ void __launch_bounds__(kTraceThreads)
k_any(DeviceScene sc, ShadowRays sh, const uint32_t* __restrict__ n_active, float tmin,
      uint32_t* __restrict__ info, uint8_t* __restrict__ occluded, TraceCounters* counters) {
    const uint32_t n = *n_active;
    TraceStats ts = {0, 0, 0};
    uint32_t rays = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float tmax = sh.tmax[i];
        if (!(tmax >= 0.0f)) continue;  // no shadow ray for this slot
        const float3 o = v3(sh.ox[i], sh.oy[i], sh.oz[i]);
        const float3 d = v3(sh.dx[i], sh.dy[i], sh.dz[i]);
        HitRecord h;
        const bool occ = traverse<true, STATS>(sc.nodes, sc.tris, sc.n_nodes, o, d, tmin, tmax, h, &ts);
        if (info && !occ) info[i] |= kInfoVisible;
        if (occluded) occluded[i] = occ ? 1 : 0;
        rays++;
    }
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->shadow_rays, (unsigned long long)rays);
    }
}

}  // namespace

void launch_closest(const LaunchCfg& lc, const DeviceScene& sc, const PathState& st,
                    const uint32_t* n_active, float tmin, float tmax, HitRecord* hits,
                    TraceCounters* counters) {
    if (lc.stats)
        k_closest<true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters);
    else
        k_closest<false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters);
    if (lc.launch_counter) ++*lc.launch_counter;
}

void launch_any(const LaunchCfg& lc, const DeviceScene& sc, const ShadowRays& sh,
                const uint32_t* n_active, float tmin, uint32_t* info, uint8_t* occluded,
                TraceCounters* counters) {
    if (lc.stats)
        k_any<true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters);
    else
        k_any<false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters);
    if (lc.launch_counter) ++*lc.launch_counter;
}

}  // namespace b2pt
