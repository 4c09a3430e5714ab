// trace_kernels.cu — K5 (closest hit) and K8 (any hit) traversal kernels.
//
// K5 replaces TraceRadiance / optixTrace(RADIANCE) (devicePrograms.cu:243-260),
// K8 replaces LightVisibility / optixTrace(SHADOW, TERMINATE_ON_FIRST_HIT)
// (:216-241) with __miss__shadow's "visible = 1" (:585-591) folded in.
// Both are persistent grid-stride kernels: the number of live rays is read
// from device memory, so the host never synchronises inside the bounce loop.
#include "wavefront.cuh"
#include "traverse.cuh"
#include "alpha.cuh"
#include "trace_wideq.cuh"

#include <cstdlib>

namespace b2pt {

namespace {

constexpr int kTraceThreads = kTraceBlock;

template <bool STATS, bool ALPHA>
__global__ void __launch_bounds__(kTraceThreads)
k_closest(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, float tmin, float tmax,
          HitRecord* __restrict__ hits, TraceCounters* counters) {
    const uint32_t n = *n_active;
    TraceStats ts = {0, 0, 0, 0, 0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float3 o = v3(st.ox[i], st.oy[i], st.oz[i]);
        const float3 d = v3(st.dx[i], st.dy[i], st.dz[i]);
        HitRecord h;
        if (ALPHA)
            traverse<false, STATS>(sc.nodes, sc.tris, sc.n_nodes, o, d, tmin, tmax, h, &ts,
                                   AlphaCutout<DeviceTexFetch>{sc, DeviceTexFetch{sc.textures}});
        else
            traverse<false, STATS>(sc.nodes, sc.tris, sc.n_nodes, o, d, tmin, tmax, h, &ts);
        reinterpret_cast<float4*>(hits)[i] =
            make_float4(__uint_as_float(h.tri), h.t, h.u, h.v);
    }
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->node_slots, (unsigned long long)ts.node_slots);
        atomicAdd(&counters->tri_slots, (unsigned long long)ts.tri_slots);
    }
}

template <bool STATS, bool ALPHA>
__global__ void __launch_bounds__(kTraceThreads)
k_any(DeviceScene sc, ShadowRays sh, const uint32_t* __restrict__ n_active, float tmin,
      uint32_t* __restrict__ info, uint8_t* __restrict__ occluded, TraceCounters* counters) {
    const uint32_t n = *n_active;
    TraceStats ts = {0, 0, 0, 0, 0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float tmax = sh.tmax[i];
        if (!(tmax >= 0.0f)) continue;  // no shadow ray for this slot
        const float3 o = v3(sh.ox[i], sh.oy[i], sh.oz[i]);
        const float3 d = v3(sh.dx[i], sh.dy[i], sh.dz[i]);
        HitRecord h;
        const bool occ =
            ALPHA ? traverse<true, STATS>(sc.nodes, sc.tris, sc.n_nodes, o, d, tmin, tmax, h, &ts,
                                          AlphaCutout<DeviceTexFetch>{sc, DeviceTexFetch{sc.textures}})
                  : traverse<true, STATS>(sc.nodes, sc.tris, sc.n_nodes, o, d, tmin, tmax, h, &ts);
        if (info && !occ) info[i] |= kInfoVisible;
        if (occluded) occluded[i] = occ ? 1 : 0;
    }
    if (STATS) {
        atomicAdd(&counters->nodes_any, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris_any, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact_any, (unsigned long long)ts.exact);
    }
}

// ---- warp-persistent variants (trace_persistent.cuh) ---------------------------
struct ClosestIo {
    PathState st;
    HitRecord* hits;
    __device__ __forceinline__ bool load(uint32_t i, float3& o, float3& d, float& tmax_out) const {
        o = v3(st.ox[i], st.oy[i], st.oz[i]);
        d = v3(st.dx[i], st.dy[i], st.dz[i]);
        tmax_out = tmax;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const HitRecord& h, bool) const {
        reinterpret_cast<float4*>(hits)[i] = make_float4(__uint_as_float(h.tri), h.t, h.u, h.v);
    }
    // deferred mode: only the triangle; t, u, v, back-face flag are evaluated by the consumer
    __device__ __forceinline__ void store_slot(uint32_t i, int slot) const {
        reinterpret_cast<float4*>(hits)[i] = make_float4(__uint_as_float(slot < 0 ? kHitMiss : (uint32_t)slot), 0.0f, 0.0f, 0.0f);
    }
    float tmax;
};
struct AnyIo {
    ShadowRays sh;
    uint32_t* info;
    uint8_t* occluded;
    __device__ __forceinline__ bool load(uint32_t i, float3& o, float3& d, float& tmax_out) const {
        tmax_out = sh.tmax[i];
        if (!(tmax_out >= 0.0f)) return false;   // no shadow ray for this slot
        o = v3(sh.ox[i], sh.oy[i], sh.oz[i]);
        d = v3(sh.dx[i], sh.dy[i], sh.dz[i]);
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const HitRecord&, bool occ) const {
        if (info && !occ) info[i] |= kInfoVisible;
        if (occluded) occluded[i] = occ ? 1 : 0;
    }
};


// K5 without the fp64 test (trace_deferred.cuh); scenes without an any-hit predicate only
template <bool STATS, int MINB = 4>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_closest_d(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, float tmin, float tmax,
            HitRecord* __restrict__ hits, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const ClosestIo io{st, hits, tmax};
    trace_closest_deferred<STATS>(sc.nodes, sc.tris, sc.n_nodes, *n_active, tmin, io, ts, pol);
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->node_slots, (unsigned long long)ts.node_slots);
        atomicAdd(&counters->tri_slots, (unsigned long long)ts.tri_slots);
    }
}

// K5 / K8 over the 4-wide nodes (trace_wide.cuh)
template <bool STATS, int MINB = 4, bool Q = false, int STEP = 0>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_closest_w(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, float tmin, float tmax,
            HitRecord* __restrict__ hits, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const ClosestIo io{st, hits, tmax};
    if constexpr (Q)
        trace_wideq<false, STATS>(sc.nodes4, sc.tris, sc.n_nodes, *n_active, tmin, io, ts, pol);
    else
        trace_wide<false, STATS, STEP>(STEP == 2 ? (const void*)sc.nodes4q : STEP == 6 ? (const void*)sc.nodes4p : (const void*)sc.nodes4, sc.tris, sc.n_nodes, sc.max_abs, *n_active, tmin, io, ts, pol);
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->node_slots, (unsigned long long)ts.node_slots);
        atomicAdd(&counters->tri_slots, (unsigned long long)ts.tri_slots);
    }
}
template <bool STATS, int MINB = 4, bool Q = false, int STEP = 0>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_any_w(DeviceScene sc, ShadowRays sh, const uint32_t* __restrict__ n_active, float tmin, uint32_t* info,
        uint8_t* occluded, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const AnyIo io{sh, info, occluded};
    if constexpr (Q)
        trace_wideq<true, STATS>(sc.nodes4, sc.tris, sc.n_nodes, *n_active, tmin, io, ts, pol);
    else
        trace_wide<true, STATS, STEP>(STEP == 2 ? (const void*)sc.nodes4q : STEP == 6 ? (const void*)sc.nodes4p : (const void*)sc.nodes4, sc.tris, sc.n_nodes, sc.max_abs, *n_active, tmin, io, ts, pol);
    if (STATS) {
        atomicAdd(&counters->nodes_any, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris_any, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact_any, (unsigned long long)ts.exact);
    }
}

// the fp64 evaluation of the hits k_closest_d found: t, u, v and the back-face flag (parity hooks;
// the renderer does this inside k_surface)
__global__ void __launch_bounds__(256)
k_finish_hits(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, HitRecord* __restrict__ hits) {
    const uint32_t n = *n_active;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = hits[i].tri;
        if (slot == kHitMiss) continue;
        hits[i] = finish_hit(sc.tris, slot, v3(st.ox[i], st.oy[i], st.oz[i]), v3(st.dx[i], st.dy[i], st.dz[i]));
    }
}

template <bool STATS, bool ALPHA, int MINB = 4>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_closest_p(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, float tmin, float tmax,
            HitRecord* __restrict__ hits, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const ClosestIo io{st, hits, tmax};
    if (ALPHA)
        trace_persistent<false, STATS>(sc.nodes, sc.tris, sc.n_nodes, *n_active, tmin, io,
                                                     AlphaCutout<DeviceTexFetch>{sc, DeviceTexFetch{sc.textures}}, ts, pol);
    else
        trace_persistent<false, STATS>(sc.nodes, sc.tris, sc.n_nodes, *n_active, tmin, io, AcceptAllHits(), ts, pol);
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->node_slots, (unsigned long long)ts.node_slots);
        atomicAdd(&counters->tri_slots, (unsigned long long)ts.tri_slots);
    }
}

template <bool STATS, bool ALPHA, int MINB = 4>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_any_p(DeviceScene sc, ShadowRays sh, const uint32_t* __restrict__ n_active, float tmin,
        uint32_t* __restrict__ info, uint8_t* __restrict__ occluded, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const AnyIo io{sh, info, occluded};
    if (ALPHA)
        trace_persistent<true, STATS>(sc.nodes, sc.tris, sc.n_nodes, *n_active, tmin, io,
                                                    AlphaCutout<DeviceTexFetch>{sc, DeviceTexFetch{sc.textures}}, ts, pol);
    else
        trace_persistent<true, STATS>(sc.nodes, sc.tris, sc.n_nodes, *n_active, tmin, io, AcceptAllHits(), ts, pol);
    if (STATS) {
        atomicAdd(&counters->nodes_any, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris_any, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact_any, (unsigned long long)ts.exact);
    }
}

// B2PT_TRAVERSAL=0 selects the first (thread-per-ray) kernels, for A/B measurements
bool use_wide();
int occ_blocks() {   // resident blocks per SM the traversal kernels are built and launched for
    static const int v = [] { const char* e = getenv("B2PT_OCC"); return e ? atoi(e) : (use_wide() ? 7 : 8); }();
    return v;
}
WarpPolicy warp_policy() {
    static const WarpPolicy v = [] {
        const char* a = getenv("B2PT_REFILL");
        const char* b = getenv("B2PT_BREAK");
        const char* q = getenv("B2PT_QBREAK");
        const char* o = getenv("B2PT_QOUT");
        int q_out = o ? atoi(o) : 1;
        q_out = q_out < 1 ? 1 : (q_out > kQMaxOut - 1 ? kQMaxOut - 1 : q_out);
        const char* sc = getenv("B2PT_SMALL_CHUNKS");
        const char* au = getenv("B2PT_ANY_UNSORTED");
        return WarpPolicy{a ? atoi(a) : 8, b ? atoi(b) : 16, q ? atoi(q) : 32, q_out, sc ? atoi(sc) : 1, au ? atoi(au) : 1};
    }();
    return v;
}
bool use_wide() {   // 4-wide nodes for K5/K8 (scenes without an any-hit predicate); B2PT_BVH4=0: binary nodes
    static const bool v = [] { const char* e = getenv("B2PT_BVH4"); return !(e && e[0] == '0'); }();
    return v;
}
// B2PT_STEP selects the node step of trace_wide (all bit-exact, A/B in profiles/r01_ab_node_step.txt):
// 6 = 96-byte nodes with the child indices in the planes' low mantissa bytes: three 256-bit loads
// (default); 5 = 128-byte nodes, three 256-bit loads + one 128-bit load of the child indices;
// 4 = four 256-bit loads (reads the 16 pad bytes too); 3 = warp-cooperative cp.async fetch (slower);
// 2 = 64-byte quantised nodes (issue-bound, slower); 1 = FMA-form slabs, seven 128-bit loads; 0 = the first step
int node_step() {
    static const int v = [] { const char* e = getenv("B2PT_STEP"); return e ? atoi(e) : 6; }();
    return v;
}
bool use_triq() {   // B2PT_TRIQ=1: warp-shared triangle queue (trace_wideq.cuh) — bit-exact, measured 2-5 % slower, kept for A/B
    static const bool v = [] { const char* e = getenv("B2PT_TRIQ"); return e && e[0] == '1'; }();
    return v;
}
bool use_deferred() {   // B2PT_DEFER=0: fp64 tests inside K5 (trace_persistent), for A/B measurements
    static const bool v = [] { const char* e = getenv("B2PT_DEFER"); return !(e && e[0] == '0'); }();
    return v;
}
bool use_persistent() {
    static const bool v = [] { const char* e = getenv("B2PT_TRAVERSAL"); return !(e && e[0] == '0'); }();
    return v;
}

}  // namespace

int trace_grid_blocks_per_sm() { return occ_blocks() > 0 ? occ_blocks() : 8; }

int trace_wide_formats() {
    if (!(use_persistent() && use_wide())) return 0;
    if (use_triq()) return 1;
    return node_step() == 2 ? 2 : node_step() == 6 ? 4 : 1;
}
bool trace_uses_wide(int any_alpha_texture) { return use_persistent() && use_wide() && any_alpha_texture == 0; }

int trace_node_bytes(const DeviceScene& sc) {
    if (use_persistent() && use_wide() && sc.any_alpha_texture == 0)
        return use_triq() ? (int)sizeof(BvhNode4) : node_step() == 2 ? (int)sizeof(BvhNode4Q) : node_step() == 6 ? (int)sizeof(BvhNode4P) : (int)sizeof(BvhNode4);
    return (int)sizeof(BvhNode2);
}

void launch_closest(const LaunchCfg& lc, const DeviceScene& sc, const PathState& st,
                    const uint32_t* n_active, float tmin, float tmax, HitRecord* hits,
                    TraceCounters* counters, bool finish) {
    const bool alpha = sc.any_alpha_texture != 0;
    if (use_persistent() && use_wide() && !alpha) {
        const bool tq = use_triq() && sc.n_tris < kQSlotLimit;
        if (lc.stats && tq)
            k_closest_w<true, 4, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (tq && occ_blocks() == 8)
            k_closest_w<false, 8, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (tq && occ_blocks() == 6)
            k_closest_w<false, 6, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (tq)
            k_closest_w<false, 4, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats && node_step() == 6)
            k_closest_w<true, 4, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 6 && occ_blocks() == 6)
            k_closest_w<false, 6, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 6 && occ_blocks() == 7)
            k_closest_w<false, 7, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 6 && occ_blocks() == 8)
            k_closest_w<false, 8, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats && node_step() == 5)
            k_closest_w<true, 4, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 5 && occ_blocks() == 6)
            k_closest_w<false, 6, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 5 && occ_blocks() == 8)
            k_closest_w<false, 8, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 5 && occ_blocks() == 7)
            k_closest_w<false, 7, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats && node_step() == 4)
            k_closest_w<true, 4, false, 4><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 4 && occ_blocks() == 6)
            k_closest_w<false, 6, false, 4><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 4 && occ_blocks() == 8)
            k_closest_w<false, 8, false, 4><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats && node_step() == 3)
            k_closest_w<true, 4, false, 3><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 3 && occ_blocks() == 6)
            k_closest_w<false, 6, false, 3><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 3 && occ_blocks() == 8)
            k_closest_w<false, 8, false, 3><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats && node_step() == 2)
            k_closest_w<true, 4, false, 2><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 2 && occ_blocks() == 6)
            k_closest_w<false, 6, false, 2><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 2 && occ_blocks() == 8)
            k_closest_w<false, 8, false, 2><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats && node_step() == 1)
            k_closest_w<true, 4, false, 1><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 1 && occ_blocks() == 6)
            k_closest_w<false, 6, false, 1><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (node_step() == 1 && occ_blocks() == 8)
            k_closest_w<false, 8, false, 1><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats)
            k_closest_w<true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 8)
            k_closest_w<false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 6)
            k_closest_w<false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else
            k_closest_w<false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        if (lc.launch_counter) ++*lc.launch_counter;
        if (finish) {
            k_finish_hits<<<lc.grid, 256, 0, lc.stream>>>(sc, st, n_active, hits);
            if (lc.launch_counter) ++*lc.launch_counter;
        }
        return;
    }
    if (use_persistent() && use_deferred() && !alpha) {
        if (lc.stats)
            k_closest_d<true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 8)
            k_closest_d<false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 6)
            k_closest_d<false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else
            k_closest_d<false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        if (lc.launch_counter) ++*lc.launch_counter;
        if (finish) {
            k_finish_hits<<<lc.grid, 256, 0, lc.stream>>>(sc, st, n_active, hits);
            if (lc.launch_counter) ++*lc.launch_counter;
        }
        return;
    }
    if (use_persistent()) {
        if (lc.stats && alpha)
            k_closest_p<true, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (lc.stats)
            k_closest_p<true, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (alpha)
            k_closest_p<false, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 8)
            k_closest_p<false, false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 6)
            k_closest_p<false, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else if (occ_blocks() == 10)
            k_closest_p<false, false, 10><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else
            k_closest_p<false, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        if (lc.launch_counter) ++*lc.launch_counter;
        return;
    }
    if (lc.stats && alpha)
        k_closest<true, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters);
    else if (lc.stats)
        k_closest<true, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters);
    else if (alpha)
        k_closest<false, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters);
    else
        k_closest<false, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters);
    if (lc.launch_counter) ++*lc.launch_counter;
}

void launch_any(const LaunchCfg& lc, const DeviceScene& sc, const ShadowRays& sh,
                const uint32_t* n_active, float tmin, uint32_t* info, uint8_t* occluded,
                TraceCounters* counters) {
    const bool alpha = sc.any_alpha_texture != 0;
    if (use_persistent() && use_wide() && !alpha) {
        const bool tq = use_triq() && sc.n_tris < kQSlotLimit;
        if (lc.stats && tq)
            k_any_w<true, 4, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (tq && occ_blocks() == 8)
            k_any_w<false, 8, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (tq && occ_blocks() == 6)
            k_any_w<false, 6, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (tq)
            k_any_w<false, 4, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats && node_step() == 6)
            k_any_w<true, 4, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 6 && occ_blocks() == 6)
            k_any_w<false, 6, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 6 && occ_blocks() == 7)
            k_any_w<false, 7, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 6 && occ_blocks() == 8)
            k_any_w<false, 8, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats && node_step() == 5)
            k_any_w<true, 4, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 5 && occ_blocks() == 6)
            k_any_w<false, 6, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 5 && occ_blocks() == 8)
            k_any_w<false, 8, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 5 && occ_blocks() == 7)
            k_any_w<false, 7, false, 5><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats && node_step() == 4)
            k_any_w<true, 4, false, 4><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 4 && occ_blocks() == 6)
            k_any_w<false, 6, false, 4><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 4 && occ_blocks() == 8)
            k_any_w<false, 8, false, 4><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats && node_step() == 3)
            k_any_w<true, 4, false, 3><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 3 && occ_blocks() == 6)
            k_any_w<false, 6, false, 3><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 3 && occ_blocks() == 8)
            k_any_w<false, 8, false, 3><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats && node_step() == 2)
            k_any_w<true, 4, false, 2><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 2 && occ_blocks() == 6)
            k_any_w<false, 6, false, 2><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 2 && occ_blocks() == 8)
            k_any_w<false, 8, false, 2><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats && node_step() == 1)
            k_any_w<true, 4, false, 1><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 1 && occ_blocks() == 6)
            k_any_w<false, 6, false, 1><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (node_step() == 1 && occ_blocks() == 8)
            k_any_w<false, 8, false, 1><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats)
            k_any_w<true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (occ_blocks() == 8)
            k_any_w<false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (occ_blocks() == 6)
            k_any_w<false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else
            k_any_w<false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        if (lc.launch_counter) ++*lc.launch_counter;
        return;
    }
    if (use_persistent()) {
        if (lc.stats && alpha)
            k_any_p<true, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats)
            k_any_p<true, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (alpha)
            k_any_p<false, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (occ_blocks() == 8)
            k_any_p<false, false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (occ_blocks() == 6)
            k_any_p<false, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (occ_blocks() == 10)
            k_any_p<false, false, 10><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else
            k_any_p<false, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        if (lc.launch_counter) ++*lc.launch_counter;
        return;
    }
    if (lc.stats && alpha)
        k_any<true, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters);
    else if (lc.stats)
        k_any<true, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters);
    else if (alpha)
        k_any<false, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters);
    else
        k_any<false, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters);
    if (lc.launch_counter) ++*lc.launch_counter;
}

}  // namespace b2pt
