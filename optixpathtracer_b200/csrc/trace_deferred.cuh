// trace_deferred.cuh — K5 for scenes without an any-hit predicate: closest hit WITHOUT evaluating the
// fp64 test in the traversal kernel.
//
// The hit contract is unchanged: the ray hits the triangle whose fp64 Moeller-Trumbore evaluation
// (tri_exact) accepts with the smallest t, ties to the lower primitive id.  What changes is where the
// fp64 arithmetic runs.  In trace_persistent() every triangle that passes the fp32 filter is evaluated
// in fp64 on the spot — 0.8 evaluations per ray, executed with 2 of 32 lanes active, 16-18 % of the
// kernel's issued instructions (ncu source view).  Here the kernel only decides
// WHICH triangle is hit:
//   * tri_classify (trace_persistent.cuh) proves most hits in fp32: barycentrics and distance clear
//     their error bounds, so tri_exact is guaranteed to accept, with t inside [t_lo, t_hi] (directed
//     rounding);
//   * a proven hit replaces the current candidate when its interval lies entirely below the
//     candidate's, and is dropped when it lies entirely above — both orderings hold for the fp64 values;
//   * only when a hit cannot be proven, or two intervals overlap (rays through a shared edge, ties),
//     the fp64 evaluation runs in the kernel, through one out-of-line routine.
// The kernel stores the triangle slot; t, u, v and the back-face flag are evaluated where they are
// consumed — K6 (k_surface), every lane with a hit — or by k_finish_hits for the parity hooks.
// Per-lane state shrinks from (hit record, best t in fp64, best primitive, fp32 bound) to
// (candidate, t_lo, t_hi), which also relieves the 64-register build.
#pragma once

#include "trace_persistent.cuh"

namespace b2pt {

// tri_classify's proven case with the interval: returns 0 / 1 / 2 like tri_classify
__device__ __forceinline__ int tri_classify_interval(float3 o, float3 d, float d_l1, float3 v0, float3 v1, float3 v2,
                                                     float tmin, float tbest, float& t_lo, float& t_hi) {
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
    const float tn_lo = __fadd_rd(T, -err_t), tn_hi = __fadd_ru(T, err_t);
    const bool inside = U > err_u && V > err_v && __fadd_ru(__fadd_ru(U, V), err_uv) < D && d_lo > 0.0f &&
                        tn_lo > __fmul_ru(tmin, d_hi) && tn_hi < __fmul_rd(tbest, d_lo);
    if (!inside || !(d_lo > 1e-30f)) return 1;
    // any enclosure of [tn_lo / d_hi, tn_hi / d_lo] will do (it only orders candidates; ties and overlaps go to
    // fp64): the approximate quotient (<= 2 ulp, MUFU.RCP + FMUL) widened by 2^-21 instead of two IEEE
    // directed-rounding divisions, which were 7 % of the kernel's issued instructions at 1-2 active lanes
    t_lo = __fmul_rd(__fdividef(tn_lo, d_hi), 0.9999995f);
    t_hi = __fmul_ru(__fdividef(tn_hi, d_lo), 1.0000005f);
    return 2;
}

struct Candidate {
    int slot;        // triangle slot or -1
    float lo, hi;    // its fp64 t lies in [lo, hi]
};

// The rare path: triangle `x` could not be proven or ordered in fp32.  Evaluates it in fp64 and, when
// needed, the current candidate too, and returns the new candidate.  One copy, out of line.
__device__ __noinline__ Candidate resolve_in_fp64(const TriRecord* __restrict__ tris, float3 o, float3 d, float tmin,
                                                  float tmax, int x, Candidate c) {
    const TriRecord* xp = tris + x;
    const float4 xa = B2PT_LDG4(&xp->a), xb = B2PT_LDG4(&xp->b), xc = B2PT_LDG4(&xp->c);
    double tx, ux, vx;
    bool bf;
    if (!tri_exact(o, d, v3(xa.x, xa.y, xa.z), v3(xa.w, xb.x, xb.y), v3(xb.z, xb.w, xc.x), tx, ux, vx, bf)) return c;
    if (!(tx > (double)tmin && tx < (double)tmax)) return c;
    const float x_lo = __double2float_rd(tx), x_hi = __double2float_ru(tx);
    bool x_wins = c.slot < 0 || x_hi < c.lo;
    if (!x_wins && !(x_lo > c.hi)) {   // still unordered: the candidate's fp64 value decides
        const TriRecord* cp = tris + c.slot;
        const float4 ca = B2PT_LDG4(&cp->a), cb = B2PT_LDG4(&cp->b), cc = B2PT_LDG4(&cp->c);
        double tc, uc, vc;
        bool bfc;
        const bool ok = tri_exact(o, d, v3(ca.x, ca.y, ca.z), v3(ca.w, cb.x, cb.y), v3(cb.z, cb.w, cc.x), tc, uc, vc, bfc);
        // (ok is guaranteed by the proof that made it the candidate; if it ever failed, x takes over)
        x_wins = !ok || tx < tc || (tx == tc && f2u_bits(xc.y) < f2u_bits(cc.y));
        if (!x_wins) {   // keep the candidate, now with its exact distance
            c.lo = __double2float_rd(tc);
            c.hi = __double2float_ru(tc);
        }
    }
    if (x_wins) {
        c.slot = x;
        c.lo = x_lo;
        c.hi = x_hi;
    }
    return c;
}

// The same decision for a triangle `x` whose fp64 evaluation the caller already holds (it accepted the hit at distance tx,
// inside the ray's range: the alpha cut-out path of the wide kernels): only the current candidate may still need its own.
__device__ __noinline__ Candidate resolve_known(const TriRecord* __restrict__ tris, float3 o, float3 d, int x, double tx,
                                                uint32_t prim_x, Candidate c) {
    const float x_lo = __double2float_rd(tx), x_hi = __double2float_ru(tx);
    bool x_wins = c.slot < 0 || x_hi < c.lo;
    if (!x_wins && !(x_lo > c.hi)) {   // unordered: the candidate's fp64 value decides, ties to the lower primitive id
        const TriRecord* cp = tris + c.slot;
        const float4 ca = B2PT_LDG4(&cp->a), cb = B2PT_LDG4(&cp->b), cc = B2PT_LDG4(&cp->c);
        double tc, uc, vc;
        bool bfc;
        const bool ok = tri_exact(o, d, v3(ca.x, ca.y, ca.z), v3(ca.w, cb.x, cb.y), v3(cb.z, cb.w, cc.x), tc, uc, vc, bfc);
        x_wins = !ok || tx < tc || (tx == tc && prim_x < f2u_bits(cc.y));
        if (!x_wins) {
            c.lo = __double2float_rd(tc);
            c.hi = __double2float_ru(tc);
        }
    }
    if (x_wins) {
        c.slot = x;
        c.lo = x_lo;
        c.hi = x_hi;
    }
    return c;
}

// Io: bool load(uint32_t i, float3& o, float3& d, float& tmax);  void store_slot(uint32_t i, int slot /* -1: miss */)
template <bool STATS, class Io>
__device__ __forceinline__ void trace_closest_deferred(const BvhNode2* __restrict__ nodes,
                                                       const TriRecord* __restrict__ tris, int n_nodes, uint32_t n_rays,
                                                       float tmin, const Io& io, TraceStats& ts, WarpPolicy pol) {
    const int kRefill = pol.refill_lanes;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lanes_lt = (1u << lane) - 1u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t chunk_sh = ray_chunk_shift(n_rays, n_warps, pol);
    uint32_t seq = 0;

    int node = kStackSentinel;
    int pend = 0;
    bool has_ray = false;
    uint32_t ri = 0;
    RaySetup r;
    r.o = r.d = r.idir = v3(0.0f);
    r.d_l1 = 0.0f;
    float tmax_f = 0.0f;
    Candidate c;          // c.hi doubles as the fp32 pruning bound while a candidate exists
    c.slot = -1;
    c.lo = 0.0f;
    c.hi = 0.0f;          // = tmax while there is no candidate
    __shared__ int s_stack[kSharedStack][kTraceBlock];
    int deep_stack[kStackSize - kSharedStack];
    int sp = 0;
    const int tid = threadIdx.x;
#define B2PT_PUSH(v) do { const int v_ = (v); if (sp < kSharedStack) s_stack[sp][tid] = v_; else deep_stack[sp - kSharedStack] = v_; sp++; } while (0)
#define B2PT_POP() (--sp, sp < kSharedStack ? s_stack[sp][tid] : deep_stack[sp - kSharedStack])

    while (true) {
        const bool idle = node == kStackSentinel;
        if (idle && has_ray) {
            io.store_slot(ri, c.slot);
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
                    tmax_f = tmax;
                    c.slot = -1;
                    c.lo = 0.0f;
                    c.hi = tmax;
                    sp = 0;
                    B2PT_PUSH(kStackSentinel);
                    node = 0;
                    pend = 0;
                } else if (i < n_rays) {
                    if (n_nodes <= 0) io.store_slot(i, -1);
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
                const BvhNode2* np = nodes + node;
                float4 a, b, cz, m;
                ldg256(&np->c0xy, a, b);
                ldg256(&np->cz, cz, m);
                float t0, t1;
                const bool h0 = slab(r, a.x, a.y, a.z, a.w, cz.x, cz.y, tmin, c.hi, t0);
                const bool h1 = slab(r, b.x, b.y, b.z, b.w, cz.z, cz.w, tmin, c.hi, t1);
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
                float t_lo = 0.0f, t_hi = 0.0f;
                const int cls = tri_classify_interval(r.o, r.d, r.d_l1, v0, v1, v2, tmin, c.hi, t_lo, t_hi);
                if (cls == 2 && (c.slot < 0 || t_hi < c.lo)) {
                    // proven hit (its t < c.hi is part of the proof) below the candidate: the new candidate
                    c.slot = slot;
                    c.lo = t_lo;
                    c.hi = t_hi;
                } else if (cls != 0) {
                    // a proven hit wholly above the candidate cannot occur (t < c.hi was proven); what is
                    // left is unproven or overlapping: fp64
                    if (STATS) ts.exact++;
                    c = resolve_in_fp64(tris, r.o, r.d, tmin, tmax_f, slot, c);
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

}  // namespace b2pt
