// trace_wide.cuh — K5 / K8 over the 4-wide nodes (BvhNode4, bvh.cuh), scenes without an any-hit predicate.
//
// Same warp-persistent loop, same deferred hit decision (trace_deferred.cuh) and same shadow-ray
// shortcut as the binary kernels; what changes is the step: one visit reads 7 x LDG.128 and tests the
// four grandchildren, so a ray takes half as many DEPENDENT fetches to reach its leaves — the binary
// kernels are bound by exactly that chain.  Children that are hit are visited
// nearest first: the nearest becomes the next node, the others go on the stack farthest first.
#pragma once

#include "trace_deferred.cuh"

namespace b2pt {

// a descent crosses at least one binary level (the greedy collapse, K3d, may open only one side) and
// pushes at most three entries; the binary tree is at most kStackSize levels deep
constexpr int kWideStackSize = 3 * kStackSize;

// slab test of slot k of a wide node (same arithmetic as slab(), traverse.cuh)
#define B2PT_SLAB4(k, comp)                                                                                   \
    {                                                                                                         \
        const float x0 = (lox.comp - r.o.x) * r.idir.x, x1 = (hix.comp - r.o.x) * r.idir.x;                   \
        const float y0 = (loy.comp - r.o.y) * r.idir.y, y1 = (hiy.comp - r.o.y) * r.idir.y;                   \
        const float z0 = (loz.comp - r.o.z) * r.idir.z, z1 = (hiz.comp - r.o.z) * r.idir.z;                   \
        const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));              \
        const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), bound));             \
        tk[k] = tn <= tf * 1.000001f ? tn : INFINITY;                                                         \
    }

// ANY = false: closest hit, deferred (Io::store_slot).  ANY = true: occlusion (Io::store(i, hit, occluded)).
template <bool ANY, bool STATS, class Io>
__device__ __forceinline__ void trace_wide(const BvhNode4* __restrict__ nodes, const TriRecord* __restrict__ tris,
                                           int n_nodes, uint32_t n_rays, float tmin, const Io& io, TraceStats& ts,
                                           WarpPolicy pol) {
    const int kRefill = pol.refill_lanes;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lanes_lt = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t chunk_sh = ray_chunk_shift(n_rays, n_warps, pol);
    uint32_t seq = 0;

    int node = kStackSentinel;
    int pend = 0;
    bool has_ray = false, occluded = false;
    uint32_t ri = 0;
    RaySetup r;
    r.o = r.d = r.idir = v3(0.0f);
    r.d_l1 = 0.0f;
    float tmax_f = 0.0f;
    Candidate c;
    c.slot = -1;
    c.lo = 0.0f;
    c.hi = 0.0f;
    __shared__ int s_stack[kSharedStack][kTraceBlock];
    int deep_stack[kWideStackSize - kSharedStack];
    int sp = 0;
    const int tid = threadIdx.x;
#define B2PT_PUSH(v) do { const int v_ = (v); if (sp < kSharedStack) s_stack[sp][tid] = v_; else deep_stack[sp - kSharedStack] = v_; sp++; } while (0)
#define B2PT_POP() (--sp, sp < kSharedStack ? s_stack[sp][tid] : deep_stack[sp - kSharedStack])

    while (true) {
        const bool idle = node == kStackSentinel;
        if (idle && has_ray) {
            if constexpr (ANY) {
                HitRecord none;
                none.tri = kHitMiss;
                none.t = none.u = none.v = 0.0f;
                io.store(ri, none, occluded);
            } else {
                io.store_slot(ri, c.slot);
            }
            has_ray = false;
        }
        const uint32_t idle_mask = __ballot_sync(kFullMask, idle);
        const uint32_t chunk_first = (((seq >> chunk_sh) * n_warps + warp) << chunk_sh);
        const bool exhausted = chunk_first >= n_rays;
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
                    tmax_f = tmax;
                    c.slot = -1;
                    c.lo = 0.0f;
                    c.hi = tmax;
                    sp = 0;
                    B2PT_PUSH(kStackSentinel);
                    node = 0;
                    pend = 0;
                } else if (i < n_rays && n_nodes <= 0) {
                    if constexpr (ANY) {
                        HitRecord none;
                        none.tri = kHitMiss;
                        none.t = none.u = none.v = 0.0f;
                        float3 o2, d2;
                        float tm2;
                        if (io.load(i, o2, d2, tm2)) io.store(i, none, false);
                    } else {
                        io.store_slot(i, -1);
                    }
                }
            }
            seq += (uint32_t)__popc(idle_mask);
            continue;
        }

        do {
            while (true) {
              if (STATS) ts.node_slots++;
              if (node >= 0 && node != kStackSentinel) {
                if (STATS) ts.nodes++;
#ifdef B2PT_DEBUG_WIDE
                if (node >= n_nodes) printf("wide: node %d of %d sp %d ray %u\n", node, n_nodes, sp, ri);
                const BvhNode4* np = nodes + (node < n_nodes ? node : 0);
#else
                const BvhNode4* np = nodes + node;
#endif
                const float4 lox = B2PT_LDG4(&np->lox), hix = B2PT_LDG4(&np->hix);
                const float4 loy = B2PT_LDG4(&np->loy), hiy = B2PT_LDG4(&np->hiy);
                const float4 loz = B2PT_LDG4(&np->loz), hiz = B2PT_LDG4(&np->hiz);
                const int4 ch = __ldg(reinterpret_cast<const int4*>(np->child));
                const float bound = c.hi;
                float tk[4];
                B2PT_SLAB4(0, x) B2PT_SLAB4(1, y) B2PT_SLAB4(2, z) B2PT_SLAB4(3, w)
                int ck[4] = {ch.x, ch.y, ch.z, ch.w};
                // sort the four (t, child) pairs by t, misses (t = inf) last: 5-comparator network
#define B2PT_CSWAP(a, b) { if (tk[b] < tk[a]) { const float tt = tk[a]; tk[a] = tk[b]; tk[b] = tt; const int cc = ck[a]; ck[a] = ck[b]; ck[b] = cc; } }
                B2PT_CSWAP(0, 1) B2PT_CSWAP(2, 3) B2PT_CSWAP(0, 2) B2PT_CSWAP(1, 3) B2PT_CSWAP(1, 2)
#undef B2PT_CSWAP
                if (tk[3] < INFINITY) B2PT_PUSH(ck[3]);
                if (tk[2] < INFINITY) B2PT_PUSH(ck[2]);
                if (tk[1] < INFINITY) B2PT_PUSH(ck[1]);
                node = tk[0] < INFINITY ? ck[0] : B2PT_POP();
                if (node < 0 && pend >= 0) {
                    pend = node;
                    node = B2PT_POP();
                }
              }
              const bool stopped = node < 0 || node == kStackSentinel;
              if ((int)__popc(__ballot_sync(kFullMask, stopped)) >= pol.break_lanes) break;
            }
            const uint32_t tris_before = ts.tris;
            while (pend < 0) {
                const int slot = ~pend;
                pend = 0;
                if (STATS) ts.tris++;
                const TriRecord* tp = tris + slot;
                const float4 ta = B2PT_LDG4(&tp->a);
                const float4 tb = B2PT_LDG4(&tp->b);
                const float tcx = __ldg(&tp->c.x);
                const float3 v0 = v3(ta.x, ta.y, ta.z), v1 = v3(ta.w, tb.x, tb.y), v2 = v3(tb.z, tb.w, tcx);
                if constexpr (ANY) {
                    const int cls = tri_classify(r.o, r.d, r.d_l1, v0, v1, v2, tmin, tmax_f);
                    bool occ = cls == 2;
                    if (cls == 1) {
                        if (STATS) ts.exact++;
                        Candidate none;
                        none.slot = -1;
                        none.lo = 0.0f;
                        none.hi = tmax_f;
                        occ = resolve_in_fp64(tris, r.o, r.d, tmin, tmax_f, slot, none).slot >= 0;
                    }
                    if (occ) {
                        occluded = true;
                        node = kStackSentinel;
                        break;
                    }
                } else {
                    float t_lo = 0.0f, t_hi = 0.0f;
                    const int cls = tri_classify_interval(r.o, r.d, r.d_l1, v0, v1, v2, tmin, c.hi, t_lo, t_hi);
                    if (cls == 2 && (c.slot < 0 || t_hi < c.lo)) {
                        c.slot = slot;
                        c.lo = t_lo;
                        c.hi = t_hi;
                    } else if (cls != 0) {
                        if (STATS) ts.exact++;
                        c = resolve_in_fp64(tris, r.o, r.d, tmin, tmax_f, slot, c);
                    }
                }
                if (node < 0) {
                    pend = node;
                    node = B2PT_POP();
                }
            }
            if (STATS) ts.tri_slots += __reduce_max_sync(kFullMask, ts.tris - tris_before);
        } while ((int)__popc(__ballot_sync(kFullMask, node == kStackSentinel)) < kRefill);
    }
#undef B2PT_PUSH
#undef B2PT_POP
}

#undef B2PT_SLAB4

}  // namespace b2pt
