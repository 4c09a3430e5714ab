// trace_wideq.cuh — K5 / K8 over the 4-wide nodes with a WARP-SHARED TRIANGLE QUEUE.
//
// trace_wide() runs the triangle code with the lanes that happen to hold a leaf: ncu's source view of that
// kernel (profiles/r01_k_closest_wide_full_summary.txt, SASS page) shows 36 % of the issued instructions in
// the triangle phase at 2-6 active lanes of 32, plus 7 % in the directed-rounding divisions behind it — a
// ray meets only ~3 leaves in ~15 node visits, so at any moment a third of the lanes hold one.
//
// Here a leaf is not bound to the lane that found it.  A lane that pops a leaf appends (triangle slot, its
// lane id) to a queue its warp keeps in shared memory (ballot + popc ranks, no atomics) and keeps
// descending; up to kQMaxOut entries per lane may be outstanding.  When the queue holds pol.q_break entries
// (or break_lanes lanes cannot continue) the warp leaves the descent and ALL 32 lanes drain the queue: lane
// k takes entry k, fetches the owner's ray with seven shuffles, runs the same fp32 classification as before
// against the owner's current pruning bound, and leaves a non-trivial outcome in shared memory for the
// owner (one shared-memory atomicOr of its lane bit into the owner's mask).  The owner then applies the
// outcomes addressed to it to its candidate — the same replace / drop / resolve-in-fp64 rule as
// trace_deferred.cuh, plus one case the queue creates: two proven hits of one ray classified against the
// same (stale) bound, the second lying wholly above the first, which is dropped.
//
// The hit contract is untouched (minimum fp64 t, ties to the lower primitive id; any-hit = existence): the
// order in which candidates meet, and how late the pruning bound tightens, cannot change either.
#pragma once

#include "trace_wide.cuh"

namespace b2pt {

constexpr int kQMaxOut = 3;               // queue entries one lane may have outstanding (pol.q_out while descending + 1 at the flush)
constexpr int kQCap = 32 * kQMaxOut;      // per warp
constexpr int kQSlotLimit = 1 << 27;      // entry = slot << 5 | lane

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
__device__ __forceinline__ void trace_wideq(const BvhNode4* __restrict__ nodes, const TriRecord* __restrict__ tris,
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
    bool has_ray = false, occluded = false;
    uint32_t ri = 0;
    float3 ro = v3(0.0f), rd = v3(0.0f), ridir = v3(0.0f);
    float tmax_f = 0.0f;
    Candidate c;
    c.slot = -1;
    c.lo = 0.0f;
    c.hi = 0.0f;
    __shared__ int s_stack[kSharedStack][kTraceBlock];
    __shared__ uint32_t s_queue[kTraceBlock / 32][kQCap];
    __shared__ float4 s_res[kTraceBlock];     // per tester lane: slot (bits), t_lo, t_hi, class (bits)
    __shared__ uint32_t s_mine[kTraceBlock];  // per owner lane: tester lanes that hold an outcome for it
    int deep_stack[kWideStackSize - kSharedStack];
    int sp = 0;
    const int tid = threadIdx.x;
    const int wbase = tid & ~31;
    uint32_t* const q = s_queue[tid >> 5];
    int qn = 0;     // entries in the warp's queue (warp-uniform)
    int nout = 0;   // of which mine
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
                const uint32_t qi = seq + (uint32_t)__popc(idle_mask & lanes_lt);
                const uint32_t i = ((((qi >> chunk_sh) * n_warps + warp) << chunk_sh)) + (qi & ((1u << chunk_sh) - 1u));
                float3 o, d;
                float tmax;
                if (i < n_rays && n_nodes > 0 && io.load(i, o, d, tmax)) {
                    const RaySetup r = ray_setup(o, d);
                    ro = r.o;
                    rd = r.d;
                    ridir = r.idir;
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
            // ---- descent: every lane with an internal node visits it; popped leaves go to the queue ----
            while (true) {
                if (STATS) ts.node_slots++;
                if (node >= 0 && node != kStackSentinel) {
                    if (STATS) ts.nodes++;
                    struct { float3 o, idir; } r = {ro, ridir};
                    const BvhNode4* np = nodes + node;
                    const float4 lox = B2PT_LDG4(&np->lox), hix = B2PT_LDG4(&np->hix);
                    const float4 loy = B2PT_LDG4(&np->loy), hiy = B2PT_LDG4(&np->hiy);
                    const float4 loz = B2PT_LDG4(&np->loz), hiz = B2PT_LDG4(&np->hiz);
                    const int4 ch = __ldg(reinterpret_cast<const int4*>(np->child));
                    const float bound = ANY ? tmax_f : c.hi;
                    float tk[4];
                    B2PT_SLAB4(0, x) B2PT_SLAB4(1, y) B2PT_SLAB4(2, z) B2PT_SLAB4(3, w)
                    int ck[4] = {ch.x, ch.y, ch.z, ch.w};
#define B2PT_CSWAP(a, b) { if (tk[b] < tk[a]) { const float tt = tk[a]; tk[a] = tk[b]; tk[b] = tt; const int cc = ck[a]; ck[a] = ck[b]; ck[b] = cc; } }
                    B2PT_CSWAP(0, 1) B2PT_CSWAP(2, 3) B2PT_CSWAP(0, 2) B2PT_CSWAP(1, 3) B2PT_CSWAP(1, 2)
#undef B2PT_CSWAP
                    if (tk[3] < INFINITY) B2PT_PUSH(ck[3]);
                    if (tk[2] < INFINITY) B2PT_PUSH(ck[2]);
                    if (tk[1] < INFINITY) B2PT_PUSH(ck[1]);
                    node = tk[0] < INFINITY ? ck[0] : B2PT_POP();
                }
                const bool push = node < 0 && nout < pol.q_out;
                const uint32_t pm = __ballot_sync(kFullMask, push);
                if (pm) {
                    if (push) {
                        q[qn + (int)__popc(pm & lanes_lt)] = ((uint32_t)(~node) << 5) | lane;
                        nout++;
                        node = B2PT_POP();
                    }
                    qn += (int)__popc(pm);
                }
                const bool stopped = node < 0 || node == kStackSentinel;
                if (qn >= pol.q_break || (int)__popc(__ballot_sync(kFullMask, stopped)) >= pol.break_lanes) break;
            }
            // ---- triangle phase: all lanes drain the queue, 32 entries per pass ----
            {   // a lane stopped by a leaf hands it in too
                const bool fl = node < 0;
                const uint32_t fm = __ballot_sync(kFullMask, fl);
                if (fl) {
                    q[qn + (int)__popc(fm & lanes_lt)] = ((uint32_t)(~node) << 5) | lane;
                    node = B2PT_POP();
                }
                qn += (int)__popc(fm);
            }
            __syncwarp();
            while (qn > 0) {
                const int base = qn > 32 ? qn - 32 : 0;
                const bool have = (int)lane < qn - base;
                const uint32_t e = have ? q[base + (int)lane] : lane;
                const int owner = (int)(e & 31u);
                const int slot = (int)(e >> 5);
                const float3 o = v3(__shfl_sync(kFullMask, ro.x, owner), __shfl_sync(kFullMask, ro.y, owner),
                                    __shfl_sync(kFullMask, ro.z, owner));
                const float3 d = v3(__shfl_sync(kFullMask, rd.x, owner), __shfl_sync(kFullMask, rd.y, owner),
                                    __shfl_sync(kFullMask, rd.z, owner));
                const float bound = __shfl_sync(kFullMask, ANY ? tmax_f : c.hi, owner);
                s_mine[tid] = 0u;
                __syncwarp();
                if (have) {
                    if (STATS) ts.tris++;
                    const TriRecord* tp = tris + slot;
                    const float4 ta = B2PT_LDG4(&tp->a);
                    const float4 tb = B2PT_LDG4(&tp->b);
                    const float tcx = __ldg(&tp->c.x);
                    const float3 v0 = v3(ta.x, ta.y, ta.z), v1 = v3(ta.w, tb.x, tb.y), v2 = v3(tb.z, tb.w, tcx);
                    const float d_l1 = fabsf(d.x) + fabsf(d.y) + fabsf(d.z);
                    float t_lo = 0.0f, t_hi = 0.0f;
                    int cls;
                    if constexpr (ANY)
                        cls = tri_classify(o, d, d_l1, v0, v1, v2, tmin, bound);
                    else
                        cls = tri_classify_interval(o, d, d_l1, v0, v1, v2, tmin, bound, t_lo, t_hi);
                    if (cls != 0) {
                        s_res[tid] = make_float4(__int_as_float(slot), t_lo, t_hi, __int_as_float(cls));
                        atomicOr(&s_mine[wbase + owner], 1u << lane);
                    }
                }
                __syncwarp();
                uint32_t m = s_mine[tid];
                while (m) {
                    const int b = __ffs((int)m) - 1;
                    m &= m - 1u;
                    const float4 rr = s_res[wbase + b];
                    const int x = __float_as_int(rr.x);
                    const int xc = __float_as_int(rr.w);
                    if constexpr (ANY) {
                        if (!occluded) {
                            bool occ = xc == 2;
                            if (!occ) {
                                if (STATS) ts.exact++;
                                Candidate none;
                                none.slot = -1;
                                none.lo = 0.0f;
                                none.hi = tmax_f;
                                occ = resolve_in_fp64(tris, ro, rd, tmin, tmax_f, x, none).slot >= 0;
                            }
                            if (occ) {
                                occluded = true;
                                node = kStackSentinel;
                            }
                        }
                    } else {
                        if (xc == 2 && (c.slot < 0 || rr.z < c.lo)) {
                            // proven hit wholly below the candidate: the new candidate
                            c.slot = x;
                            c.lo = rr.y;
                            c.hi = rr.z;
                        } else if (xc == 2 && rr.y > c.hi) {
                            // proven against a bound that an earlier outcome of this pass has since lowered:
                            // wholly above the candidate, so farther in fp64 too
                        } else {
                            if (STATS) ts.exact++;
                            c = resolve_in_fp64(tris, ro, rd, tmin, tmax_f, x, c);
                        }
                    }
                }
                __syncwarp();
                if (STATS) ts.tri_slots++;
                qn = base;
            }
            nout = 0;
        } while ((int)__popc(__ballot_sync(kFullMask, node == kStackSentinel)) < kRefill);
    }
#undef B2PT_PUSH
#undef B2PT_POP
}

#undef B2PT_SLAB4

}  // namespace b2pt
