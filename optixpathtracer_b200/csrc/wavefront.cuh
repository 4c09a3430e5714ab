// wavefront.cuh — data layout of the wavefront integrator and kernel launchers.
//
// The reference is an OptiX megakernel: one thread per pixel runs the whole
// path loop (devicePrograms.cu:625-664) with its state in a 64-byte
// RadianceRayData reached through two payload registers (RayData.h:5-51), and
// closest-hit does all shading plus a nested shadow trace (:343-514).
// Here the path loop is unrolled in time: every bounce is a short sequence of
// kernels over structure-of-arrays path state, so each kernel is coherent
// (traversal only traverses, each BSDF kernel only evaluates one BSDF) and
// every global access of the state is coalesced.
//
// Per bounce:   K5 closest-hit  ->  K6 surface + classify + light pick
//            -> K8 shadow any-hit -> K7a..d shade (NEE add, sample, next ray,
//               termination, compaction into the other state buffer)
// Per wave:     K4 raygen before the first bounce, K9 accumulate after the last.
#pragma once

#include "scene.cuh"

namespace b2pt {

// SoA path state; one slot per in-flight path.  Two copies (ping-pong): the
// shade kernels write survivors densely into the other copy (A3 in SURVEY §8).
struct PathState {
    float *ox, *oy, *oz;      // next ray origin      (RadianceRayData.nextOrigin)
    float *dx, *dy, *dz;      // next ray direction   (.nextDirection)
    float *bx, *by, *bz;      // throughput           (.beta)
    float *rx, *ry, *rz;      // radiance so far      (.radiance)
    uint32_t* id;             // pixel + W*H*local_sample: where the result goes
    uint32_t* seed;           // LCG state (.randomSeed); unused with Philox
};
constexpr int kPathStateWords = 14;

// Per-hit record written by K6 and read by K7 (Surface.h:5-39, minus what can
// be recomputed from it): 56 bytes, SoA.
struct SurfaceState {
    float *px, *py, *pz;      // surface.position
    float *gx, *gy, *gz;      // surface.gNormal (world)
    float *nx, *ny, *nz;      // surface.sNormal (world); the shading frame is rebuilt from it
    float *ar, *ag, *ab;      // surface.albedo
    float* rough;             // surface.roughness
    uint32_t* info;           // bits 0-7 light index, 8 has-light, 9 conductor, 10 visible (set by K8)
};
constexpr int kSurfaceStateWords = 14;
constexpr uint32_t kInfoHasLight = 1u << 8;
constexpr uint32_t kInfoConductor = 1u << 9;
constexpr uint32_t kInfoVisible = 1u << 10;

// Shadow rays, one per slot (LightVisibility, devicePrograms.cu:216-241).
struct ShadowRays {
    float *ox, *oy, *oz, *dx, *dy, *dz;
    float* tmax;              // < 0: no shadow ray for this slot
    // Light-ordered work list (optional, the renderer's K6 fills it; the parity hooks leave it null): the slots
    // that DO hold a shadow ray, appended to bucket (light index & 3), so that the rays a warp of K8 traces
    // together aim at the same light and the empty slots are never visited.  order[k * capacity + j] = slot of
    // the j-th ray of bucket k, bucket_n[k] = rays in bucket k.  The answer of a ray does not depend on where it
    // sits in the list (K8 writes by slot).
    uint32_t* order = nullptr;      // [kShadowBuckets][capacity]
    uint32_t* bucket_n = nullptr;   // [kShadowBuckets], zeroed with the wave's counters
    uint32_t capacity = 0;
};
constexpr int kShadowBuckets = 4;
constexpr int kShadowWords = 7;

struct TraceCounters {
    unsigned long long nodes, tris, exact, radiance_rays, shadow_rays, nan_samples;
    unsigned long long nodes_any, tris_any, exact_any;
    unsigned long long node_slots, tri_slots;   // K5 stats build: lane slots of the node / triangle loops
    unsigned long long shadow_culled;   // shadow rays K6 did not emit because f(wo, wl) is exactly zero
    unsigned long long walk_f_hits, walk_s_hits;   // layered walk engine: hits with f walks (5 tasks each) / sample walks
};
// kClsEngine: the walk-engine kernel of the layered BSDF (K7d-e) on its own; reported inside the shade class too
enum KernelClass : int { kClsClosest = 0, kClsAny = 1, kClsSurface = 2, kClsShade = 3, kClsOther = 4, kClsEngine = 5, kClsCount = 6 };

// ---- launchers (wavefront.cu / trace_kernels.cu) ---------------------------
struct LaunchCfg {
    cudaStream_t stream;
    int grid;            // persistent grid: multiple of the SM count
    bool stats;
    uint64_t* launch_counter;
};

// resident traversal blocks per SM the launchers size their grids for (B2PT_OCC overrides)
int trace_grid_blocks_per_sm();
// bytes of the node record K5/K8 read for this scene (64: binary nodes, 128: 4-wide nodes)
int trace_node_bytes(const DeviceScene& sc);


// K5: closest hit for rays [0, *n_active) of `st`; n_active lives on the device.
void launch_closest(const LaunchCfg& lc, const DeviceScene& sc, const PathState& st,
                    const uint32_t* n_active, float tmin, float tmax, HitRecord* hits,
                    TraceCounters* counters, bool finish = true);
// finish = false: the caller evaluates t, u, v and the back-face flag of the stored triangle itself
// (finish_hit, traverse.cuh) — k_surface does; true adds k_finish_hits after the deferred K5.
// K8: any hit for shadow rays [0, *n_active); sets kInfoVisible in info[] (if
// info != nullptr) and/or occluded[] (if non-null).
void launch_any(const LaunchCfg& lc, const DeviceScene& sc, const ShadowRays& sh,
                const uint32_t* n_active, float tmin, uint32_t* info, uint8_t* occluded,
                TraceCounters* counters);

}  // namespace b2pt
