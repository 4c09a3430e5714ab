// trace_wide.cuh — K5 / K8 over the 4-wide nodes (BvhNode4P / BvhNode4, bvh.cuh): the kernels of every scene.
//
// Same warp-persistent loop, same deferred hit decision (trace_deferred.cuh) and same shadow-ray
// shortcut as the binary kernels; what changes is the step: one visit reads three 256-bit loads and tests
// the four grandchildren, so a ray takes half as many DEPENDENT fetches to reach its leaves — the binary
// kernels are bound by exactly that chain.  Children that are hit are visited
// nearest first: the nearest becomes the next node, the others go on the stack farthest first.
//
// Any-hit predicate (alpha cut-out, alpha.cuh): triangles of meshes WITHOUT an albedo texture take the
// fp32-proof path unchanged; a triangle of a textured mesh is evaluated in fp64 on the spot (filter ->
// tri_exact -> predicate at its u, v) and only an accepted hit enters the candidate logic, so the result is
// the reference's (closest ACCEPTED hit, ties to the lower primitive id) while untextured geometry of a
// textured scene runs at the speed of an untextured scene.
#pragma once

#include "trace_deferred.cuh"

namespace b2pt {

// kWideStackSize (bvh.cuh): a visit pushes at most three entries; the build refuses trees whose measured wide
// depth would overflow it

// Slab test of slot k of a wide node with one FMA per plane: t = plane * (1/d) - o * (1/d), the second product rounded once per
// ray.  Its rounding error, <= 2^-24 |o| in position units, is absolute, not relative like the errors
// tf * 1.000001f covers; the boxes' outward padding (2^-17 of the scene's largest coordinate M, bvh_build.cu) absorbs it
// with a factor 8 to spare while |o| <= 16 M.  Rays that start farther out (a distant camera) carry the bound
// as explicit slack, eslack = 2^-22 max_i |o_i / d_i| (0 otherwise) — correct for them too, only less tight.
#define B2PT_SLAB4F(k, comp)                                                                                  \
    {                                                                                                         \
        const float x0 = fmaf(lox.comp, r.idir.x, noid.x), x1 = fmaf(hix.comp, r.idir.x, noid.x);             \
        const float y0 = fmaf(loy.comp, r.idir.y, noid.y), y1 = fmaf(hiy.comp, r.idir.y, noid.y);             \
        const float z0 = fmaf(loz.comp, r.idir.z, noid.z), z1 = fmaf(hiz.comp, r.idir.z, noid.z);             \
        const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));              \
        const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), bound));             \
        tk[k] = tn <= fmaf(tf, 1.000001f, eslack) ? tn : INFINITY;                                            \
    }

// A triangle of a cut-out mesh that passed the fp32 filter: exact evaluation, predicate at its (u, v), and — closest-hit
// rays — the candidate update with the fp64 distance in hand.  One copy, out of line.
template <class Accept>
__device__ __noinline__ Candidate cutout_triangle(const TriRecord* __restrict__ tris, const Accept& accept, float3 o, float3 d,
                                                  float tmin, float tmax, int slot, uint32_t mesh, uint32_t prim, Candidate c,
                                                  bool any, bool& accepted) {
    const TriRecord* tp = tris + slot;
    const float4 ta = B2PT_LDG4(&tp->a), tb = B2PT_LDG4(&tp->b);
    const float tcx = __ldg(&tp->c.x);
    double t, u, v;
    bool bf;
    accepted = tri_exact(o, d, v3(ta.x, ta.y, ta.z), v3(ta.w, tb.x, tb.y), v3(tb.z, tb.w, tcx), t, u, v, bf) &&
               t > (double)tmin && t < (double)tmax && accept((uint32_t)slot, mesh, (float)u, (float)v);
    if (accepted && !any) c = resolve_known(tris, o, d, slot, t, prim, c);
    return c;
}

// ANY = false: closest hit, deferred (Io::store_slot).  ANY = true: occlusion (Io::store(i, hit, occluded)).
// The node step: FMA-form slabs (above), branch-free pushes while the stack stays inside its shared-memory
// levels, an unconditional read of the stack top instead of a divergent pop.
// STEP = 6 (default): the 96-byte node BvhNode4P (child indices in the low mantissa bytes of the x / y planes),
// three 256-bit loads (LDG.E.256, new on sm_100) per visit.  STEP = 5: the 128-byte node BvhNode4, three 256-bit
// loads for the boxes and one 128-bit load for the child indices.  (The earlier steps — seven 128-bit loads,
// a warp-cooperative cp.async fetch, 64-byte quantised nodes, a warp-shared triangle queue — all measured
// slower, profiles/r01_ab_node_step.txt, and are no longer built.)
// SS = stack levels kept in shared memory ([level][thread]); deeper levels spill to local memory.
template <bool ANY, bool STATS, int STEP, int SS, class Io, class Accept>
__device__ __forceinline__ void trace_wide(const void* __restrict__ nodes_v, const TriRecord* __restrict__ tris,
                                           int n_nodes, float max_abs, uint32_t n_rays, float tmin, const Io& io,
                                           const Accept& accept, TraceStats& ts, WarpPolicy pol) {
    static_assert(STEP == 5 || STEP == 6, "node steps 0-4 are no longer built");
    constexpr bool kPredicate = !std::is_same<Accept, AcceptAllHits>::value;
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
    float3 noid = v3(0.0f);   // -(o / d), per component
    float eslack = 0.0f;
    float tmax_f = 0.0f;
    Candidate c;
    c.slot = -1;
    c.lo = 0.0f;
    c.hi = 0.0f;
    __shared__ int s_stack[SS][kTraceBlock];
    int deep_stack[kWideStackSize - SS];
    int sp = 0;
    const int tid = threadIdx.x;
#define B2PT_PUSH(v) do { const int v_ = (v); if (sp < SS) s_stack[sp][tid] = v_; else deep_stack[sp - SS] = v_; sp++; } while (0)
#define B2PT_POP() (--sp, sp < SS ? s_stack[sp][tid] : deep_stack[sp - SS])

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
                    noid = v3(-(o.x * r.idir.x), -(o.y * r.idir.y), -(o.z * r.idir.z));
                    const float far_o = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
                    eslack = far_o <= 16.0f * max_abs
                                 ? 0.0f
                                 : 2.3841858e-7f * (far_o + max_abs) *
                                       fmaxf(fmaxf(fabsf(r.idir.x), fabsf(r.idir.y)), fabsf(r.idir.z));
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
                const float bound = c.hi;
                float tk[4];
                int4 ch;
                if constexpr (STEP == 6) {
                    const BvhNode4P* np = static_cast<const BvhNode4P*>(nodes_v) + node;
                    float4 lox, hix, loy, hiy, loz, hiz;
                    ldg256(&np->lox, lox, hix);
                    ldg256(&np->loy, loy, hiy);
                    ldg256(&np->loz, loz, hiz);
                    // child k = the low bytes of lox[k], hix[k], loy[k], hiy[k]
#define B2PT_CHILD(c) ((int)__byte_perm(__byte_perm(__float_as_uint(lox.c), __float_as_uint(hix.c), 0x0040u),   \
                                        __byte_perm(__float_as_uint(loy.c), __float_as_uint(hiy.c), 0x0040u), 0x5410u))
                    ch = make_int4(B2PT_CHILD(x), B2PT_CHILD(y), B2PT_CHILD(z), B2PT_CHILD(w));
#undef B2PT_CHILD
                    B2PT_SLAB4F(0, x) B2PT_SLAB4F(1, y) B2PT_SLAB4F(2, z) B2PT_SLAB4F(3, w)
                } else {
                    const BvhNode4* np = static_cast<const BvhNode4*>(nodes_v) + node;
                    float4 lox, hix, loy, hiy, loz, hiz;
                    ldg256(&np->lox, lox, hix);
                    ldg256(&np->loy, loy, hiy);
                    ldg256(&np->loz, loz, hiz);
                    ch = __ldg(reinterpret_cast<const int4*>(np->child));
                    B2PT_SLAB4F(0, x) B2PT_SLAB4F(1, y) B2PT_SLAB4F(2, z) B2PT_SLAB4F(3, w)
                }
                int ck[4] = {ch.x, ch.y, ch.z, ch.w};
                // sort the four (t, child) pairs by t, misses (t = inf) last: 5-comparator network
#define B2PT_CSWAP(a, b) { if (tk[b] < tk[a]) { const float tt = tk[a]; tk[a] = tk[b]; tk[b] = tt; const int cc = ck[a]; ck[a] = ck[b]; ck[b] = cc; } }
                if (!(ANY && pol.any_unsorted)) {   // occlusion rays may skip the ordering (A/B: B2PT_ANY_UNSORTED)
                    B2PT_CSWAP(0, 1) B2PT_CSWAP(2, 3) B2PT_CSWAP(0, 2) B2PT_CSWAP(1, 3) B2PT_CSWAP(1, 2)
                }
#undef B2PT_CSWAP
                if (sp + 3 <= SS) {
                    // each store lands at the current top and the top only moves for a hit, so the entries are
                    // contiguous, farthest first; a store for a miss is overwritten by the next one
                    s_stack[sp][tid] = ck[3]; sp += tk[3] < INFINITY ? 1 : 0;
                    s_stack[sp][tid] = ck[2]; sp += tk[2] < INFINITY ? 1 : 0;
                    s_stack[sp][tid] = ck[1]; sp += tk[1] < INFINITY ? 1 : 0;
                    const int top = s_stack[sp - 1][tid];   // sp >= 1: the sentinel sits at the bottom
                    const bool go = tk[0] < INFINITY;
                    node = go ? ck[0] : top;
                    sp -= go ? 0 : 1;
                } else {
                    if (tk[3] < INFINITY) B2PT_PUSH(ck[3]);
                    if (tk[2] < INFINITY) B2PT_PUSH(ck[2]);
                    if (tk[1] < INFINITY) B2PT_PUSH(ck[1]);
                    node = tk[0] < INFINITY ? ck[0] : B2PT_POP();
                }
                if (node < 0 && pend >= 0) {
                    pend = node;
                    node = B2PT_POP();
                }
              }
              const bool stopped = node < 0 || node == kStackSentinel;
              if ((int)__popc(__ballot_sync(kFullMask, stopped)) >= (ANY ? pol.break_lanes_any : pol.break_lanes)) break;
            }
            const uint32_t tris_before = ts.tris;
            while (pend < 0) {
                // a leaf names a run of triangle records (bvh.cuh: first | (count - 1) << 28); the run is walked one
                // record per trip, the rest of it stays parked in `pend`
                const uint32_t leaf = ~(uint32_t)pend;
                const int slot = (int)(leaf & kLeafSlotMask);
                pend = leaf > kLeafSlotMask ? (int)~(leaf + 1u - (1u << kLeafCountShift)) : 0;
                if (STATS) ts.tris++;
                // (reading the 48-byte record as one 256-bit + one 128-bit load, chosen by the slot's parity, was
                // measured 1-2 % slower than these three loads: profiles/r01_ab_node_step.txt)
                const TriRecord* tp = tris + slot;
                const float4 ta = B2PT_LDG4(&tp->a);
                const float4 tb = B2PT_LDG4(&tp->b);
                float tcx;
                uint32_t mesh = 0, prim = 0;
                bool textured = false;
                if constexpr (kPredicate) {
                    // the cut-out flag rides in the record (TriRecord::c.w, bvh.cuh): no materials[mesh] load per test
                    const float4 tc = B2PT_LDG4(&tp->c);
                    tcx = tc.x;
                    prim = f2u_bits(tc.y);
                    mesh = f2u_bits(tc.z);
                    textured = f2u_bits(tc.w) != 0u;
                } else {
                    tcx = __ldg(&tp->c.x);
                }
                const float3 v0 = v3(ta.x, ta.y, ta.z), v1 = v3(ta.w, tb.x, tb.y), v2 = v3(tb.z, tb.w, tcx);
                if (textured) {
                    // the predicate needs the exact u, v: filter -> fp64 -> texel, as the binary kernels do — out of line
                    // (cutout_triangle), so that the registers of the fp64 test and of the texture fetch are not held
                    // through the node loop: untextured geometry of a textured scene runs the untextured kernel's code
                    if (tri_filter(r.o, r.d, r.d_l1, v0, v1, v2, tmin, ANY ? tmax_f : c.hi)) {
                        if (STATS) ts.exact++;
                        bool accepted = false;
                        c = cutout_triangle(tris, accept, r.o, r.d, tmin, tmax_f, slot, mesh, prim, c, ANY, accepted);
                        if constexpr (ANY) {
                            if (accepted) {
                                occluded = true;
                                node = kStackSentinel;
                                pend = 0;
                                break;
                            }
                        }
                    }
                } else if constexpr (ANY) {
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
                        pend = 0;
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
                if (pend == 0 && node < 0) {
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

#undef B2PT_SLAB4F

}  // namespace b2pt
