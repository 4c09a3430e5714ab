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

// kWideStackSize (bvh.cuh): a visit pushes at most three entries; the build refuses trees whose measured wide
// depth would overflow it

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

// The same test with one FMA per plane: t = plane * (1/d) - o * (1/d), the second product rounded once per
// ray.  Its rounding error, <= 2^-24 |o| in position units, is absolute, not relative like the errors
// tf * 1.000001f covers; the boxes' outward padding (2^-18 of the scene's largest coordinate M, K3) absorbs it
// with a factor 4 to spare while |o| <= 16 M.  Rays that start farther out (a distant camera) carry the bound
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

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// ANY = false: closest hit, deferred (Io::store_slot).  ANY = true: occlusion (Io::store(i, hit, occluded)).
// DIET: the node step with fewer instructions — FMA-form slabs (above), branch-free pushes while the stack
// stays inside its shared-memory levels, an unconditional read of the stack top instead of a divergent pop.
// Same hits (both forms of the slab test are conservative); the traversal order can differ where two entry
// distances compare differently.
// STEP = 2: DIET over the 64-byte quantised nodes (BvhNode4Q): 4 loads per visit instead of 7.
// STEP = 4: DIET over the fp32 nodes read with four 256-bit loads (LDG.E.256, new on sm_100) instead of seven
// 128-bit ones.  STEP = 5 (default): three 256-bit loads for the boxes and one 128-bit load for the child
// indices — the pad bytes of the node are not read.  ncu: a divergent 256-bit load is not free, the L1TEX
// data pipe still moves it as more than one wavefront (8.3 M wavefronts per SM against 9.0 M for the 128-bit
// form on the same launch), but the launch is 6-8 % shorter.
// STEP = 6: the 96-byte node BvhNode4P (child indices in the low mantissa bytes of the x / y planes): three
// 256-bit loads per visit.
// STEP = 3: DIET over the fp32 nodes, fetched by the warp TOGETHER.  A lane's seven 16-byte loads of its own
// node cost seven L1TEX wavefronts — divergent lanes, one wavefront per lane per instruction, 16 of the
// pipe's 128 bytes used; ncu of the fp32 step: L1TEX 88 % busy, long-scoreboard 5.1 of 10.6 cycles per
// issue (profiles/r01_k_closest_step1_vs_step2.txt).  Here eight lanes copy one lane's node — a whole
// 128-byte line, ONE wavefront — with cp.async (global -> shared, no registers), four nodes per
// instruction, eight instructions for the 32 nodes of a warp; each lane then reads its node back with
// seven conflict-free LDS.128 (the 16-byte pieces are XOR-swizzled by the owner's lane).
template <bool ANY, bool STATS, int STEP, class Io>
__device__ __forceinline__ void trace_wide(const void* __restrict__ nodes_v, const TriRecord* __restrict__ tris,
                                           int n_nodes, float max_abs, uint32_t n_rays, float tmin, const Io& io,
                                           TraceStats& ts, WarpPolicy pol) {
    constexpr bool DIET = STEP >= 1;
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
    float3 noid = v3(0.0f);   // DIET: -(o / d), per component
    float eslack = 0.0f;
    float tmax_f = 0.0f;
    Candidate c;
    c.slot = -1;
    c.lo = 0.0f;
    c.hi = 0.0f;
    __shared__ int s_stack[kSharedStack][kTraceBlock];
    __shared__ float4 s_node[STEP == 3 ? kTraceBlock : 1][8];   // STEP 3: the node each lane is about to visit
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
                    if constexpr (DIET) {
                        noid = v3(-(o.x * r.idir.x), -(o.y * r.idir.y), -(o.z * r.idir.z));
                        const float far_o = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
                        eslack = far_o <= 16.0f * max_abs
                                     ? 0.0f
                                     : 2.3841858e-7f * (far_o + max_abs) *
                                           fmaxf(fmaxf(fabsf(r.idir.x), fabsf(r.idir.y)), fabsf(r.idir.z));
                    }
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
              if constexpr (STEP == 3) {
                  const uint32_t act = __ballot_sync(kFullMask, node >= 0 && node != kStackSentinel);
                  const int piece = (int)(lane & 7u), grp = (int)(lane >> 3);
                  const int wb = tid & ~31;
#pragma unroll
                  for (int rr = 0; rr < 8; rr++) {
                      if ((act >> (4 * rr)) & 0xfu) {   // warp-uniform: some lane of 4rr .. 4rr+3 visits a node
                          const int owner = 4 * rr + grp;
                          const int n_o = __shfl_sync(kFullMask, node, owner);
                          if (((act >> owner) & 1u) && piece < 7)
                              cp_async_16(&s_node[wb + owner][piece ^ (owner & 7)],
                                          static_cast<const char*>(nodes_v) + (size_t)n_o * sizeof(BvhNode4) + piece * 16);
                      }
                  }
                  cp_async_wait_all();
                  __syncwarp();
              }
              if (node >= 0 && node != kStackSentinel) {
                if (STATS) ts.nodes++;
                const float bound = c.hi;
                float tk[4];
                int4 ch;
                if constexpr (STEP == 2) {
                    const BvhNode4Q* np = static_cast<const BvhNode4Q*>(nodes_v) + node;
                    const float4 A = B2PT_LDG4(&np->ox);
                    const uint4 B = __ldg(reinterpret_cast<const uint4*>(&np->qlox));
                    const uint4 C = __ldg(reinterpret_cast<const uint4*>(&np->qloz));
                    ch = __ldg(reinterpret_cast<const int4*>(np->child));
                    const uint32_t ex = __float_as_uint(A.w);
                    // scale * (1/d): exact (power of two); (origin - o) * (1/d) with the ray's own product -(o/d)
                    const float sx = __uint_as_float((ex & 0xffu) << 23) * r.idir.x;
                    const float sy = __uint_as_float((ex & 0xff00u) << 15) * r.idir.y;
                    const float sz = __uint_as_float((ex & 0xff0000u) << 7) * r.idir.z;
                    const float bx = fmaf(A.x, r.idir.x, noid.x), by = fmaf(A.y, r.idir.y, noid.y),
                                bz = fmaf(A.z, r.idir.z, noid.z);
                    // byte k of w as a float: 0x4B0000qq is 2^23 + q exactly
#define B2PT_Q(w, k) (__uint_as_float(__byte_perm((w), 0x4B000000u, 0x7650u + (k))) - 8388608.0f)
#define B2PT_SLAB4Q(k, cid)                                                                                   \
    {                                                                                                         \
        const float x0 = fmaf(B2PT_Q(B.x, k), sx, bx), x1 = fmaf(B2PT_Q(B.y, k), sx, bx);                     \
        const float y0 = fmaf(B2PT_Q(B.z, k), sy, by), y1 = fmaf(B2PT_Q(B.w, k), sy, by);                     \
        const float z0 = fmaf(B2PT_Q(C.x, k), sz, bz), z1 = fmaf(B2PT_Q(C.y, k), sz, bz);                     \
        const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));              \
        const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), bound));             \
        tk[k] = (cid) != kEmptyChild && tn <= fmaf(tf, 1.000001f, eslack) ? tn : INFINITY;                    \
    }
                    B2PT_SLAB4Q(0, ch.x) B2PT_SLAB4Q(1, ch.y) B2PT_SLAB4Q(2, ch.z) B2PT_SLAB4Q(3, ch.w)
#undef B2PT_SLAB4Q
#undef B2PT_Q
                } else if constexpr (STEP == 3) {
                    const float4* sn = s_node[tid];
                    const int sw = (int)(lane & 7u);
                    const float4 lox = sn[0 ^ sw], hix = sn[1 ^ sw], loy = sn[2 ^ sw], hiy = sn[3 ^ sw];
                    const float4 loz = sn[4 ^ sw], hiz = sn[5 ^ sw];
                    const float4 chf = sn[6 ^ sw];
                    ch = make_int4(__float_as_int(chf.x), __float_as_int(chf.y), __float_as_int(chf.z), __float_as_int(chf.w));
                    B2PT_SLAB4F(0, x) B2PT_SLAB4F(1, y) B2PT_SLAB4F(2, z) B2PT_SLAB4F(3, w)
                } else if constexpr (STEP == 6) {
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
                } else if constexpr (STEP == 4 || STEP == 5) {
                    const BvhNode4* np = static_cast<const BvhNode4*>(nodes_v) + node;
                    float4 lox, hix, loy, hiy, loz, hiz;
                    ldg256(&np->lox, lox, hix);
                    ldg256(&np->loy, loy, hiy);
                    ldg256(&np->loz, loz, hiz);
                    if constexpr (STEP == 4) {
                        float4 chf, padf;
                        ldg256(np->child, chf, padf);
                        ch = make_int4(__float_as_int(chf.x), __float_as_int(chf.y), __float_as_int(chf.z), __float_as_int(chf.w));
                    } else {
                        ch = __ldg(reinterpret_cast<const int4*>(np->child));
                    }
                    B2PT_SLAB4F(0, x) B2PT_SLAB4F(1, y) B2PT_SLAB4F(2, z) B2PT_SLAB4F(3, w)
                } else {
                    const BvhNode4* np = static_cast<const BvhNode4*>(nodes_v) + node;
                    const float4 lox = B2PT_LDG4(&np->lox), hix = B2PT_LDG4(&np->hix);
                    const float4 loy = B2PT_LDG4(&np->loy), hiy = B2PT_LDG4(&np->hiy);
                    const float4 loz = B2PT_LDG4(&np->loz), hiz = B2PT_LDG4(&np->hiz);
                    ch = __ldg(reinterpret_cast<const int4*>(np->child));
                    if constexpr (DIET) {
                        B2PT_SLAB4F(0, x) B2PT_SLAB4F(1, y) B2PT_SLAB4F(2, z) B2PT_SLAB4F(3, w)
                    } else {
                        B2PT_SLAB4(0, x) B2PT_SLAB4(1, y) B2PT_SLAB4(2, z) B2PT_SLAB4(3, w)
                    }
                }
                int ck[4] = {ch.x, ch.y, ch.z, ch.w};
                // sort the four (t, child) pairs by t, misses (t = inf) last: 5-comparator network
#define B2PT_CSWAP(a, b) { if (tk[b] < tk[a]) { const float tt = tk[a]; tk[a] = tk[b]; tk[b] = tt; const int cc = ck[a]; ck[a] = ck[b]; ck[b] = cc; } }
                if (!(ANY && pol.any_unsorted)) {   // occlusion rays may skip the ordering (A/B: B2PT_ANY_UNSORTED)
                    B2PT_CSWAP(0, 1) B2PT_CSWAP(2, 3) B2PT_CSWAP(0, 2) B2PT_CSWAP(1, 3) B2PT_CSWAP(1, 2)
                }
#undef B2PT_CSWAP
                if (DIET && sp + 3 <= kSharedStack) {
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
              if ((int)__popc(__ballot_sync(kFullMask, stopped)) >= pol.break_lanes) break;
            }
            const uint32_t tris_before = ts.tris;
            while (pend < 0) {
                const int slot = ~pend;
                pend = 0;
                if (STATS) ts.tris++;
                // (reading the 48-byte record as one 256-bit + one 128-bit load, chosen by the slot's parity, was
                // measured 1-2 % slower than these three loads: profiles/r01_ab_node_step.txt)
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
#undef B2PT_SLAB4F

}  // namespace b2pt
