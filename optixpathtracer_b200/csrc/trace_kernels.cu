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
#include "trace_wide.cuh"

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
    // position in the light-ordered work list -> slot (ShadowRays::order).  The bucket sizes are re-read from
    // memory (L1) at each use — a few times per ray — instead of living in registers through the traversal loop.
    __device__ __forceinline__ uint32_t slot_of(uint32_t q) const {
        if (!sh.order) return q;
        uint32_t k = 0, base = 0;
#pragma unroll
        for (int b = 0; b < kShadowBuckets - 1; b++) {
            const uint32_t nb = __ldg(sh.bucket_n + b);
            if (q >= base + nb && k == (uint32_t)b) {
                base += nb;
                k = (uint32_t)b + 1u;
            }
        }
        return sh.order[(size_t)k * sh.capacity + (q - base)];
    }
    __device__ __forceinline__ uint32_t count(uint32_t n_slots) const {
        if (!sh.order) return n_slots;
        uint32_t n = 0;
#pragma unroll
        for (int b = 0; b < kShadowBuckets; b++) n += __ldg(sh.bucket_n + b);
        return n;
    }
    __device__ __forceinline__ bool load(uint32_t q, float3& o, float3& d, float& tmax_out) const {
        const uint32_t i = slot_of(q);
        tmax_out = sh.tmax[i];
        if (!(tmax_out >= 0.0f)) return false;   // no shadow ray for this slot
        o = v3(sh.ox[i], sh.oy[i], sh.oz[i]);
        d = v3(sh.dx[i], sh.dy[i], sh.dz[i]);
        return true;
    }
    __device__ __forceinline__ void store(uint32_t q, const HitRecord&, bool occ) const {
        const uint32_t i = slot_of(q);
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

// K5 / K8 over the 4-wide nodes (trace_wide.cuh).  ALPHA: the scene has an alpha cut-out texture (any-hit
// predicate inside the traversal); SS: stack levels in shared memory
template <bool STATS, int MINB, int STEP, bool ALPHA, int SS = kSharedStack>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_closest_w(DeviceScene sc, PathState st, const uint32_t* __restrict__ n_active, float tmin, float tmax,
            HitRecord* __restrict__ hits, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const ClosestIo io{st, hits, tmax};
    const void* nodes = STEP == 6 ? (const void*)sc.nodes4p : (const void*)sc.nodes4;
    if constexpr (ALPHA)
        trace_wide<false, STATS, STEP, SS>(nodes, sc.tris, sc.n_nodes, sc.max_abs, *n_active, tmin, io,
                                           AlphaCutout<DeviceTexFetch>{sc, DeviceTexFetch{sc.textures}}, ts, pol);
    else
        trace_wide<false, STATS, STEP, SS>(nodes, sc.tris, sc.n_nodes, sc.max_abs, *n_active, tmin, io, AcceptAllHits(), ts, pol);
    if (STATS) {
        atomicAdd(&counters->nodes, (unsigned long long)ts.nodes);
        atomicAdd(&counters->tris, (unsigned long long)ts.tris);
        atomicAdd(&counters->exact, (unsigned long long)ts.exact);
        atomicAdd(&counters->node_slots, (unsigned long long)ts.node_slots);
        atomicAdd(&counters->tri_slots, (unsigned long long)ts.tri_slots);
    }
}
template <bool STATS, int MINB, int STEP, bool ALPHA, int SS = kSharedStack>
__global__ void __launch_bounds__(kTraceThreads, MINB)
k_any_w(DeviceScene sc, ShadowRays sh, const uint32_t* __restrict__ n_active, float tmin, uint32_t* info,
        uint8_t* occluded, TraceCounters* counters, WarpPolicy pol) {
    TraceStats ts = {0, 0, 0, 0, 0};
    const AnyIo io{sh, info, occluded};
    const void* nodes = STEP == 6 ? (const void*)sc.nodes4p : (const void*)sc.nodes4;
    const uint32_t n_rays = io.count(*n_active);   // the light-ordered list holds only the slots with a ray
    if constexpr (ALPHA)
        trace_wide<true, STATS, STEP, SS>(nodes, sc.tris, sc.n_nodes, sc.max_abs, n_rays, tmin, io,
                                          AlphaCutout<DeviceTexFetch>{sc, DeviceTexFetch{sc.textures}}, ts, pol);
    else
        trace_wide<true, STATS, STEP, SS>(nodes, sc.tris, sc.n_nodes, sc.max_abs, n_rays, tmin, io, AcceptAllHits(), ts, pol);
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


// ---- switches (A/B measurements; the defaults are what the bench runs) ----------------------------
// B2PT_TRAVERSAL=0 selects the first (thread-per-ray) kernels; B2PT_BVH4=0 the binary-node warp-persistent ones
// (B2PT_DEFER=0: with the fp64 tests inside K5).  Every variant is bit-exact (tests/test_variants_gpu.py).
bool use_wide() {
    static const bool v = [] { const char* e = getenv("B2PT_BVH4"); return !(e && e[0] == '0'); }();
    return v;
}
// B2PT_STEP: 6 = 96-byte nodes, three 256-bit loads (default); 5 = 128-byte nodes, three 256-bit loads + one
// 128-bit load of the child indices (profiles/r01_ab_node_step.txt)
int node_step() {
    static const int v = [] { const char* e = getenv("B2PT_STEP"); return e && atoi(e) == 5 ? 5 : 6; }();
    return v;
}
bool use_deferred() {
    static const bool v = [] { const char* e = getenv("B2PT_DEFER"); return !(e && e[0] == '0'); }();
    return v;
}
bool use_persistent() {
    static const bool v = [] { const char* e = getenv("B2PT_TRAVERSAL"); return !(e && e[0] == '0'); }();
    return v;
}
// B2PT_SSTACK = 16 (default) | 24 | 32: traversal-stack levels of the wide kernels kept in shared memory
int shared_stack_levels() {
    static const int v = [] {
        const char* e = getenv("B2PT_SSTACK");
        const int n = e ? atoi(e) : kSharedStack;
        return n == 24 || n == 32 ? n : kSharedStack;
    }();
    return v;
}
// blocks per SM of the alpha-predicate instantiation (B2PT_ALPHA_OCC = 6 | 7; A/B: profiles/r02_ab_cutout.txt)
int alpha_occ() {
    static const int v = [] { const char* e = getenv("B2PT_ALPHA_OCC"); return e ? atoi(e) : 7; }();
    return v;
}
constexpr int kWideOcc = 7;   // resident blocks per SM the wide kernels are built for (A/B: profiles/r01_ab_traversal_occupancy.txt)
int occ_blocks() {            // ... and the binary ones: B2PT_OCC = 4 | 6 | 8 (default) | 10
    static const int v = [] { const char* e = getenv("B2PT_OCC"); return e ? atoi(e) : 8; }();
    return v;
}
WarpPolicy warp_policy() {
    static const WarpPolicy v = [] {
        const char* a = getenv("B2PT_REFILL");
        const char* b = getenv("B2PT_BREAK");
        const char* sc = getenv("B2PT_SMALL_CHUNKS");
        const char* au = getenv("B2PT_ANY_UNSORTED");
        const char* ba = getenv("B2PT_BREAK_ANY");
        // break thresholds 16 (K5) / 20 (K8): profiles/r02_ab_warp_policy.txt
        return WarpPolicy{a ? atoi(a) : 8, b ? atoi(b) : 16, sc ? atoi(sc) : 1, au ? atoi(au) : 1, ba ? atoi(ba) : (b ? atoi(b) : 20)};
    }();
    return v;
}

}  // namespace

int trace_grid_blocks_per_sm() { return use_persistent() && use_wide() ? kWideOcc : (occ_blocks() > 0 ? occ_blocks() : 8); }

// wide-node formats the build must keep (bit 0 BvhNode4, bit 2 BvhNode4P)
int trace_wide_formats() {
    if (!(use_persistent() && use_wide())) return 0;
    return node_step() == 6 ? 4 : 1;
}
// every scene goes through the wide kernels (the alpha cut-out predicate runs inside them)
bool trace_uses_wide(int /*any_alpha_texture*/) { return use_persistent() && use_wide(); }

int trace_node_bytes(const DeviceScene&) {
    if (use_persistent() && use_wide()) return node_step() == 6 ? (int)sizeof(BvhNode4P) : (int)sizeof(BvhNode4);
    return (int)sizeof(BvhNode2);
}

#define B2PT_WIDE_ARGS_C <<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy())
#define B2PT_WIDE_ARGS_A <<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy())
// the stats build and the alpha build hold more registers: 4 / 6 blocks per SM
#define B2PT_WIDE_LAUNCH(KERNEL, ARGS)                                                         \
    do {                                                                                       \
        const int ss = shared_stack_levels();                                                  \
        const bool s6 = node_step() == 6;                                                      \
        if (lc.stats) {                                                                        \
            if (alpha) { if (s6) KERNEL<true, 4, 6, true> ARGS; else KERNEL<true, 4, 5, true> ARGS; }        \
            else       { if (s6) KERNEL<true, 4, 6, false> ARGS; else KERNEL<true, 4, 5, false> ARGS; }      \
        } else if (alpha) {                                                                    \
            if (!s6) KERNEL<false, 6, 5, true> ARGS;                                           \
            else if (alpha_occ() == 6) KERNEL<false, 6, 6, true> ARGS;                         \
            else KERNEL<false, kWideOcc, 6, true> ARGS;                                        \
        } else if (!s6) {                                                                      \
            KERNEL<false, kWideOcc, 5, false> ARGS;                                            \
        } else if (ss == 24) {                                                                 \
            KERNEL<false, kWideOcc, 6, false, 24> ARGS;                                        \
        } else if (ss == 32) {                                                                 \
            KERNEL<false, kWideOcc, 6, false, 32> ARGS;                                        \
        } else {                                                                               \
            KERNEL<false, kWideOcc, 6, false> ARGS;                                            \
        }                                                                                      \
    } while (0)

void launch_closest(const LaunchCfg& lc, const DeviceScene& sc, const PathState& st,
                    const uint32_t* n_active, float tmin, float tmax, HitRecord* hits,
                    TraceCounters* counters, bool finish) {
    const bool alpha = sc.any_alpha_texture != 0;
    if (use_persistent() && use_wide()) {
        B2PT_WIDE_LAUNCH(k_closest_w, B2PT_WIDE_ARGS_C);
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
        else if (occ_blocks() == 6)
            k_closest_d<false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else
            k_closest_d<false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
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
        else if (occ_blocks() == 6)
            k_closest_p<false, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
        else
            k_closest_p<false, false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, st, n_active, tmin, tmax, hits, counters, warp_policy());
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

void launch_any(const LaunchCfg& lc, const DeviceScene& sc, const ShadowRays& sh_in,
                const uint32_t* n_active, float tmin, uint32_t* info, uint8_t* occluded,
                TraceCounters* counters) {
    const bool alpha = sc.any_alpha_texture != 0;
    ShadowRays sh = sh_in;
    if (use_persistent() && use_wide()) {
        B2PT_WIDE_LAUNCH(k_any_w, B2PT_WIDE_ARGS_A);
        if (lc.launch_counter) ++*lc.launch_counter;
        return;
    }
    sh.order = nullptr;   // the A/B fallback kernels scan the slots (tmax < 0 marks the empty ones)
    if (use_persistent()) {
        if (lc.stats && alpha)
            k_any_p<true, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (lc.stats)
            k_any_p<true, false><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (alpha)
            k_any_p<false, true><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else if (occ_blocks() == 6)
            k_any_p<false, false, 6><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
        else
            k_any_p<false, false, 8><<<lc.grid, kTraceThreads, 0, lc.stream>>>(sc, sh, n_active, tmin, info, occluded, counters, warp_policy());
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
