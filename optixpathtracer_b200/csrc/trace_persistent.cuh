// trace_persistent.cuh — warp-persistent BVH traversal (the K5/K8 kernel bodies).
//
// Same hit contract and the same per-triangle / per-box arithmetic as
// traverse() in traverse.cuh (which stays the single-thread statement the CPU
// test tier runs); what changes is how a WARP is kept busy — the first kernel
// of this round ran at 6-13 active lanes of 32 (profiles/r01_k_closest_full_
// summary.csv):
//
//   * rays are not bound to threads: a warp owns an interleaved sequence of
//     64-ray chunks and a lane that finishes its ray takes the next one as soon
//     as kRefill lanes are idle (warp-uniform cursor + ballot/popc ranks; no
//     atomics, the assignment is deterministic);
//   * while-while: all lanes descend internal nodes together, then all lanes
//     test their triangles together; a lane that reaches a leaf early postpones
//     it and keeps descending, so the fp64-backed triangle code runs with many
//     lanes active instead of one or two.
//
// Traversal order differs from traverse(), the result cannot: the closest hit
// is the minimum fp64 t with ties broken by primitive id, and any-hit is an
// existence test.
#pragma once

#include <type_traits>

#include "traverse.cuh"

namespace b2pt {

// tri_filter (traverse.cuh) with one more answer.  0: the fp64 test is guaranteed to reject (or to land
// outside (tmin, tbest]); 1: undecided, run it; 2: it is guaranteed to ACCEPT with t inside
// (tmin, tbest) — the fp32 barycentrics and distance clear their error bounds on the inside, evaluated
// with directed rounding.  The margins K * (...) bound |fp32 value - real value|; the fp64 evaluation
// differs from the real value by ~1e-16 of the same magnitudes, far inside them.
__device__ __forceinline__ int tri_classify(float3 o, float3 d, float d_l1, float3 v0, float3 v1, float3 v2,
                                            float tmin, float tbest) {
    const float K = 1.9073486e-6f;
    const float3 e1 = v1 - v0, e2 = v2 - v0, tv = o - v0;
    const float3 p = cross(d, e2);
    const float det = dot(e1, p);
    const float a1 = fabsf(e1.x) + fabsf(e1.y) + fabsf(e1.z);
    const float a2 = fabsf(e2.x) + fabsf(e2.y) + fabsf(e2.z);
    const float at = fabsf(tv.x) + fabsf(tv.y) + fabsf(tv.z);
    const float err_det = K * a1 * a2 * d_l1;
    const float D = fabsf(det);
    if (!(D > err_det)) return 1;  // grazing / degenerate: let fp64 decide
    const float sgn = det < 0.0f ? -1.0f : 1.0f;
    const float err_u = K * at * a2 * d_l1;
    const float U = dot(tv, p) * sgn;
    if (U < -err_u) return 0;
    const float3 q = cross(tv, e1);
    const float err_v = K * at * a1 * d_l1;
    const float V = dot(d, q) * sgn;
    if (V < -err_v) return 0;
    const float err_uv = err_det + err_u + err_v;
    if (U + V > D + err_uv) return 0;
    const float err_t = K * at * a1 * a2;
    const float T = dot(e2, q) * sgn;
    if (T + err_t < tmin * (D - err_det)) return 0;
    if (T - err_t > tbest * (D + err_det)) return 0;
    const float d_lo = __fadd_rd(D, -err_det), d_hi = __fadd_ru(D, err_det);
    const bool inside = U > err_u && V > err_v && __fadd_ru(__fadd_ru(U, V), err_uv) < D && d_lo > 0.0f &&
                        __fadd_rd(T, -err_t) > __fmul_ru(tmin, d_hi) && __fadd_ru(T, err_t) < __fmul_rd(tbest, d_lo);
    return inside ? 2 : 1;
}

// 256-bit read-only load (sm_100: LDG.E.256): two float4 of a 32-byte-aligned record in ONE L1TEX wavefront
__device__ __forceinline__ void ldg256(const void* p, float4& a, float4& b) {
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
        : "l"(p));
}

constexpr uint32_t kFullMask = 0xffffffffu;
constexpr uint32_t kRayChunkShift = 6;   // a warp takes its rays in chunks of 64 ...
constexpr int kSharedStack = 16;   // stack levels kept in shared memory
constexpr int kTraceBlock = 128;   // threads per block of the traversal kernels

// Io: bool load(uint32_t i, float3& o, float3& d, float& tmax)   (false: slot carries no ray)
//     void store(uint32_t i, const HitRecord& h, bool occluded)
struct WarpPolicy {
    int refill_lanes;   // take new rays when this many lanes are idle
    int break_lanes;    // leave the descent when this many lanes cannot descend (hold two leaves / are done)
    int small_chunks;   // 1: launches with fewer than 64 rays per warp use smaller chunks (ray_chunk_shift)
    int any_unsorted;   // 1: occlusion rays visit the hit children of a wide node in slot order, not nearest first
    int break_lanes_any;   // break_lanes of the wide occlusion kernel (its leaves end rays: it pays to wait for more of them)
};

// ... unless the launch is small.  The late bounces of a wave hold a few thousand rays (c2: 0.06 % of the paths
// reach bounce 16); in 64-ray chunks they would queue behind each other in a handful of warps while the
// other 3 500 warps of the grid exit at once, and the launch lasts 2-3 ray lifetimes of a lone warp
// (~120 us measured, whatever the ray count).  With fewer than 64 rays per warp the chunk shrinks to the
// largest power of two <= rays / warps, down to one ray per warp: every ray starts at once.
__device__ __forceinline__ uint32_t ray_chunk_shift(uint32_t n_rays, uint32_t n_warps, const WarpPolicy& pol) {
    const uint32_t per = n_rays / n_warps;
    if (!pol.small_chunks || per >= (1u << kRayChunkShift)) return kRayChunkShift;
    return per <= 1u ? 0u : 31u - (uint32_t)__clz((int)per);
}

template <bool ANY, bool STATS, class Accept, class Io>
__device__ __forceinline__ void trace_persistent(const BvhNode2* __restrict__ nodes,
                                                 const TriRecord* __restrict__ tris, int n_nodes,
                                                 uint32_t n_rays, float tmin, const Io& io,
                                                 const Accept& accept, TraceStats& ts, WarpPolicy pol) {
    const int kRefill = pol.refill_lanes;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lanes_lt = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t chunk_sh = ray_chunk_shift(n_rays, n_warps, pol);
    uint32_t seq = 0;   // warp-uniform: next position of this warp's ray sequence

    // per-lane ray state
    int node = kStackSentinel;   // sentinel = idle
    int pend = 0;                // postponed leaf (< 0) or none
    bool has_ray = false, occluded = false;
    uint32_t ri = 0;
    RaySetup r;
    r.o = r.d = r.idir = v3(0.0f);
    r.d_l1 = 0.0f;
    float best32 = 0.0f;
    double best_t = 0.0, tmax_d = 0.0;
    const double tmin_d = (double)tmin;
    uint32_t best_prim = 0xFFFFFFFFu;
    HitRecord hit;
    hit.tri = kHitMiss;
    hit.t = hit.u = hit.v = 0.0f;
    // traversal stack: the first kSharedStack levels live in shared memory ([level][thread]: every
    // lane has its own bank whatever its depth), deeper levels spill to a local array
    __shared__ int s_stack[kSharedStack][kTraceBlock];
    int deep_stack[kStackSize - kSharedStack];
    int sp = 0;
    const int tid = threadIdx.x;
#define B2PT_PUSH(v) do { const int v_ = (v); if (sp < kSharedStack) s_stack[sp][tid] = v_; else deep_stack[sp - kSharedStack] = v_; sp++; } while (0)
#define B2PT_POP() (--sp, sp < kSharedStack ? s_stack[sp][tid] : deep_stack[sp - kSharedStack])

    while (true) {
        // ---- finished lanes publish; idle lanes take new rays -----------------
        const bool idle = node == kStackSentinel;
        if (idle && has_ray) {
            io.store(ri, hit, occluded);
            has_ray = false;
        }
        const uint32_t idle_mask = __ballot_sync(kFullMask, idle);
        const uint32_t chunk_first = (((seq >> chunk_sh) * n_warps + warp) << chunk_sh);
        const bool exhausted = chunk_first >= n_rays;   // chunks only grow: nothing left for this warp
        if (exhausted && idle_mask == kFullMask) break;
        if (!exhausted && (int)__popc(idle_mask) >= kRefill) {
            if (idle) {
                const uint32_t q = seq + (uint32_t)__popc(idle_mask & lanes_lt);
                const uint32_t i = ((((q >> chunk_sh) * n_warps + warp) << chunk_sh)) + (q & ((1u << chunk_sh) - 1u));
                float3 o, d;
                float tmax;
                if (i < n_rays && n_nodes > 0 && io.load(i, o, d, tmax)) {
                    r = ray_setup(o, d);
                    ri = i;
                    has_ray = true;
                    occluded = false;
                    tmax_d = (double)tmax;
                    best_t = tmax_d;
                    best32 = tmax;
                    best_prim = 0xFFFFFFFFu;
                    hit.tri = kHitMiss;
                    hit.t = hit.u = hit.v = 0.0f;
                    sp = 0;
                    B2PT_PUSH(kStackSentinel);
                    node = 0;
                    pend = 0;
                } else if (i < n_rays) {   // empty scene or empty slot: publish the miss right away
                    HitRecord miss;
                    miss.tri = kHitMiss;
                    miss.t = miss.u = miss.v = 0.0f;
                    if (n_nodes <= 0) io.store(i, miss, false);
                }
            }
            seq += (uint32_t)__popc(idle_mask);
            continue;   // re-evaluate: slots without rays leave lanes idle
        }

        // ---- traverse until enough lanes ran dry -------------------------------
        do {
            // the warp descends together; a lane stops with a second leaf in `node`, or done, and
            // the warp leaves the descent once break_lanes lanes have stopped
            while (true) {
              if (STATS) ts.node_slots++;
              if (node >= 0 && node != kStackSentinel) {
                if (STATS) ts.nodes++;
                const BvhNode2* np = nodes + node;
                float4 a, b, c, m;   // the 64-byte node as two 256-bit loads: half the L1TEX wavefronts
                ldg256(&np->c0xy, a, b);
                ldg256(&np->cz, c, m);
                float t0, t1;
                const bool h0 = slab(r, a.x, a.y, a.z, a.w, c.x, c.y, tmin, best32, t0);
                const bool h1 = slab(r, b.x, b.y, b.z, b.w, c.z, c.w, tmin, best32, t1);
                const int c0 = (int)f2u_bits(m.x), c1 = (int)f2u_bits(m.y);
                if (h0 && h1) {
                    const bool swap = t1 < t0;
                    node = swap ? c1 : c0;
                    B2PT_PUSH(swap ? c0 : c1);
                } else if (h0) {
                    node = c0;
                } else if (h1) {
                    node = c1;
                } else {
                    node = B2PT_POP();
                }
                if (node < 0 && pend >= 0) {   // postpone the first leaf, keep descending
                    pend = node;
                    node = B2PT_POP();
                }
              }
              const bool stopped = node < 0 || node == kStackSentinel;
              if ((int)__popc(__ballot_sync(kFullMask, stopped)) >= pol.break_lanes) break;
            }
            // all lanes test their postponed triangles
            const uint32_t tris_before = ts.tris;
            while (pend < 0) {
                const int slot = ~pend;
                pend = 0;
                if (STATS) ts.tris++;
                const TriRecord* tp = tris + slot;
                const float4 ta = B2PT_LDG4(&tp->a);
                const float4 tb = B2PT_LDG4(&tp->b);
                const float4 tc = B2PT_LDG4(&tp->c);
                const float3 v0 = v3(ta.x, ta.y, ta.z), v1 = v3(ta.w, tb.x, tb.y), v2 = v3(tb.z, tb.w, tc.x);
                // shadow rays without an any-hit predicate: a triangle the fp32 test PROVES to be hit inside
                // (tmin, tmax) occludes without the fp64 test (closest hits still need its t, u, v)
                constexpr bool kProven = ANY && std::is_same<Accept, AcceptAllHits>::value;
                const int cls = kProven ? tri_classify(r.o, r.d, r.d_l1, v0, v1, v2, tmin, best32)
                                        : (tri_filter(r.o, r.d, r.d_l1, v0, v1, v2, tmin, best32) ? 1 : 0);
                if (kProven && cls == 2) {
                    occluded = true;
                    node = kStackSentinel;
                    break;
                }
                if (cls != 0) {
                    if (STATS) ts.exact++;
                    double t, u, v;
                    bool bf;
                    if (tri_exact(r.o, r.d, v0, v1, v2, t, u, v, bf) && t > tmin_d && t < tmax_d) {
                        if (ANY) {
                            if (accept((uint32_t)slot, f2u_bits(tc.z), (float)u, (float)v)) {
                                occluded = true;
                                node = kStackSentinel;   // done: drop the rest of the stack
                                break;
                            }
                        } else {
                            const uint32_t prim = f2u_bits(tc.y);
                            if ((t < best_t || (t == best_t && prim < best_prim)) &&
                                accept((uint32_t)slot, f2u_bits(tc.z), (float)u, (float)v)) {
                                best_t = t;
                                best_prim = prim;
                                float tf = (float)t;   // round up: the fp32 bound never cuts below best_t
                                if ((double)tf < t) tf = nextafterf(tf, INFINITY);
                                best32 = tf;
                                hit.tri = (uint32_t)slot | (bf ? kHitBackface : 0u);
                                hit.t = (float)t;
                                hit.u = (float)u;
                                hit.v = (float)v;
                            }
                        }
                    }
                }
                if (node < 0) {   // the descent stopped on a second leaf: take it next
                    pend = node;
                    node = B2PT_POP();
                }
            }
            if (STATS) ts.tri_slots += __reduce_max_sync(kFullMask, ts.tris - tris_before);   // summed over lanes: 32 x warp trips
        } while ((int)__popc(__ballot_sync(kFullMask, node == kStackSentinel)) < kRefill);
    }
}

#undef B2PT_PUSH
#undef B2PT_POP

}  // namespace b2pt
