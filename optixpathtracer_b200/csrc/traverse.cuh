// traverse.cuh — closest-hit / any-hit BVH traversal, per-thread logic (K5, K8).
//
// Replaces optixTrace over the OptiX GAS (devicePrograms.cu:224-238 shadow
// rays with TERMINATE_ON_FIRST_HIT, :248-259 radiance rays; any-hit programs
// :549-561 are compiled in only for alpha-textured meshes — see wavefront.cu).
//
// Hit contract ("exact hits"): a ray hits a triangle iff the
// fp64 Moeller-Trumbore test below (fixed operation order, no FMA contraction)
// says so on the fp32 world-space vertices, with tmin < t < tmax; the closest
// hit is the minimum fp64 t, ties broken by the lower global primitive id.
// The fp32 test in front of it is only a conservative FILTER with explicit
// rounding-error bounds, and the fp32 slab test is conservative too (padded
// boxes + relative slack), so the result equals a brute-force fp64 search over
// all triangles bit for bit — which is exactly what oracle/ checks.
#pragma once

#include "bvh.cuh"

namespace b2pt {

#if defined(__CUDA_ARCH__)
#define B2PT_DMUL(a, b) __dmul_rn((a), (b))
#define B2PT_DADD(a, b) __dadd_rn((a), (b))
#define B2PT_DSUB(a, b) __dsub_rn((a), (b))
#define B2PT_DDIV(a, b) __ddiv_rn((a), (b))
#else  // host test harness: compiled with -ffp-contract=off
#define B2PT_DMUL(a, b) ((a) * (b))
#define B2PT_DADD(a, b) ((a) + (b))
#define B2PT_DSUB(a, b) ((a) - (b))
#define B2PT_DDIV(a, b) ((a) / (b))
#endif

struct TraceStats {
    uint32_t nodes, tris, exact;
    uint32_t node_slots, tri_slots;   // persistent kernels: lane slots (32 per warp trip) of the node / triangle loops
};

// fp64 Moeller-Trumbore, the deciding test.  Operation order is part of the
// contract (oracle/exact_tri.h::exact_triangle, used by oracle/oracle_trace.cpp, mirrors it).
B2PT_HD bool tri_exact(float3 of, float3 df, float3 v0f, float3 v1f, float3 v2f,
                       double& t, double& u, double& v, bool& backface) {
    const double ox = of.x, oy = of.y, oz = of.z;
    const double dx = df.x, dy = df.y, dz = df.z;
    const double ax = v0f.x, ay = v0f.y, az = v0f.z;
    const double e1x = B2PT_DSUB((double)v1f.x, ax), e1y = B2PT_DSUB((double)v1f.y, ay),
                 e1z = B2PT_DSUB((double)v1f.z, az);
    const double e2x = B2PT_DSUB((double)v2f.x, ax), e2y = B2PT_DSUB((double)v2f.y, ay),
                 e2z = B2PT_DSUB((double)v2f.z, az);
    // p = d x e2
    const double px = B2PT_DSUB(B2PT_DMUL(dy, e2z), B2PT_DMUL(dz, e2y));
    const double py = B2PT_DSUB(B2PT_DMUL(dz, e2x), B2PT_DMUL(dx, e2z));
    const double pz = B2PT_DSUB(B2PT_DMUL(dx, e2y), B2PT_DMUL(dy, e2x));
    const double det = B2PT_DADD(B2PT_DADD(B2PT_DMUL(e1x, px), B2PT_DMUL(e1y, py)), B2PT_DMUL(e1z, pz));
    if (det == 0.0) return false;
    const double inv = B2PT_DDIV(1.0, det);
    const double tx = B2PT_DSUB(ox, ax), ty = B2PT_DSUB(oy, ay), tz = B2PT_DSUB(oz, az);
    u = B2PT_DMUL(B2PT_DADD(B2PT_DADD(B2PT_DMUL(tx, px), B2PT_DMUL(ty, py)), B2PT_DMUL(tz, pz)), inv);
    if (!(u >= 0.0) || u > 1.0) return false;
    // q = tvec x e1
    const double qx = B2PT_DSUB(B2PT_DMUL(ty, e1z), B2PT_DMUL(tz, e1y));
    const double qy = B2PT_DSUB(B2PT_DMUL(tz, e1x), B2PT_DMUL(tx, e1z));
    const double qz = B2PT_DSUB(B2PT_DMUL(tx, e1y), B2PT_DMUL(ty, e1x));
    v = B2PT_DMUL(B2PT_DADD(B2PT_DADD(B2PT_DMUL(dx, qx), B2PT_DMUL(dy, qy)), B2PT_DMUL(dz, qz)), inv);
    if (!(v >= 0.0) || B2PT_DADD(u, v) > 1.0) return false;
    t = B2PT_DMUL(B2PT_DADD(B2PT_DADD(B2PT_DMUL(e2x, qx), B2PT_DMUL(e2y, qy)), B2PT_DMUL(e2z, qz)), inv);
    // front face = counter-clockwise seen from the ray origin (OptiX convention
    // behind optixIsTriangleBackFaceHit, devicePrograms.cu:349):
    // det = e1.(d x e2) = -d.(e1 x e2)  ->  back face iff det < 0.
    backface = det < 0.0;
    return true;
}

// fp32 pre-test with explicit error bounds: returns false only when the fp64
// test is guaranteed to reject (or to land outside (tmin, tbest)).
B2PT_HD bool tri_filter(float3 o, float3 d, float d_l1, float3 v0, float3 v1, float3 v2,
                        float tmin, float tbest) {
    const float K = 1.9073486e-6f;  // 32 * 2^-24: covers the ~13 roundings of each sum
    const float3 e1 = v1 - v0, e2 = v2 - v0, tv = o - v0;
    const float3 p = cross(d, e2);
    const float det = dot(e1, p);
    const float a1 = fabsf(e1.x) + fabsf(e1.y) + fabsf(e1.z);
    const float a2 = fabsf(e2.x) + fabsf(e2.y) + fabsf(e2.z);
    const float at = fabsf(tv.x) + fabsf(tv.y) + fabsf(tv.z);
    const float err_det = K * a1 * a2 * d_l1;
    const float D = fabsf(det);
    if (!(D > err_det)) return true;  // grazing / degenerate: let fp64 decide
    const float sgn = det < 0.0f ? -1.0f : 1.0f;
    const float err_u = K * at * a2 * d_l1;
    const float U = dot(tv, p) * sgn;
    if (U < -err_u) return false;
    const float3 q = cross(tv, e1);
    const float err_v = K * at * a1 * d_l1;
    const float V = dot(d, q) * sgn;
    if (V < -err_v) return false;
    if (U + V > D + (err_det + err_u + err_v)) return false;
    const float err_t = K * at * a1 * a2;
    const float T = dot(e2, q) * sgn;
    if (T + err_t < tmin * (D - err_det)) return false;
    if (T - err_t > tbest * (D + err_det)) return false;
    return true;
}

// t, u, v and the back-face flag of a known hit: the fp64 evaluation of triangle `slot` (the
// deferred K5 stores only the slot, trace_deferred.cuh)
B2PT_HD HitRecord finish_hit(const TriRecord* __restrict__ tris, uint32_t slot, float3 o, float3 d) {
    const TriRecord* tp = tris + slot;
    const float4 ta = B2PT_LDG4(&tp->a), tb = B2PT_LDG4(&tp->b), tc = B2PT_LDG4(&tp->c);
    double t, u, v;
    bool bf = false;
    HitRecord h;
    if (tri_exact(o, d, v3(ta.x, ta.y, ta.z), v3(ta.w, tb.x, tb.y), v3(tb.z, tb.w, tc.x), t, u, v, bf)) {
        h.tri = slot | (bf ? kHitBackface : 0u);
        h.t = (float)t;
        h.u = (float)u;
        h.v = (float)v;
    } else {   // cannot happen for a slot the traversal proved hit; a miss is the safe reading
        h.tri = kHitMiss;
        h.t = h.u = h.v = 0.0f;
    }
    return h;
}

struct RaySetup {
    float3 o, d, idir;
    float d_l1;
};
B2PT_HD RaySetup ray_setup(float3 o, float3 d) {
    RaySetup r;
    r.o = o;
    r.d = d;
    const float tiny = 1e-20f;
    r.idir.x = 1.0f / (fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x));
    r.idir.y = 1.0f / (fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y));
    r.idir.z = 1.0f / (fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z));
    r.d_l1 = fabsf(d.x) + fabsf(d.y) + fabsf(d.z);
    return r;
}

// Conservative slab test of one child box; tn = entry distance.
B2PT_HD bool slab(const RaySetup& r, float lox, float hix, float loy, float hiy, float loz, float hiz,
                  float tmin, float tbest, float& tn) {
    const float x0 = (lox - r.o.x) * r.idir.x, x1 = (hix - r.o.x) * r.idir.x;
    const float y0 = (loy - r.o.y) * r.idir.y, y1 = (hiy - r.o.y) * r.idir.y;
    const float z0 = (loz - r.o.z) * r.idir.z, z1 = (hiz - r.o.z) * r.idir.z;
    tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tbest));
    return tn <= tf * 1.000001f;
}

// kStackSize / kWideStackSize: bvh.cuh (the build checks the measured depth of the tree against them)
constexpr int kStackSentinel = 0x7fffffff;


// The any-hit program as a predicate on a candidate hit (A9): the default
// accepts everything and compiles to nothing; scenes with albedo textures use
// AlphaCutout (alpha.cuh).
struct AcceptAllHits {
    B2PT_HD bool textured(uint32_t /*mesh*/) const { return false; }
    B2PT_HD bool operator()(uint32_t /*slot*/, uint32_t /*mesh*/, float /*u*/, float /*v*/) const { return true; }
};

// ANY = false: closest hit into `hit`; returns hit != miss.
// ANY = true : returns true as soon as one triangle is hit in (tmin, tmax).
template <bool ANY, bool STATS, class Accept = AcceptAllHits>
B2PT_HD bool traverse(const BvhNode2* __restrict__ nodes, const TriRecord* __restrict__ tris,
                      int n_nodes, float3 o, float3 d, float tmin, float tmax, HitRecord& hit,
                      TraceStats* stats, const Accept& accept = Accept()) {
    hit.tri = kHitMiss;
    hit.t = hit.u = hit.v = 0.0f;
    if (n_nodes <= 0) return false;
    const RaySetup r = ray_setup(o, d);
    const double tmin_d = (double)tmin, tmax_d = (double)tmax;
    double best_t = tmax_d;
    float best32 = tmax;  // >= best_t always
    uint32_t best_prim = 0xFFFFFFFFu;

    int stack[kStackSize];
    int sp = 0;
    stack[sp++] = kStackSentinel;
    int node = 0;
    while (node != kStackSentinel) {
        if (node >= 0) {
            if (STATS) stats->nodes++;
            const BvhNode2* np = nodes + node;
            const float4 a = B2PT_LDG4(&np->c0xy);
            const float4 b = B2PT_LDG4(&np->c1xy);
            const float4 c = B2PT_LDG4(&np->cz);
            const float4 m = B2PT_LDG4(&np->child0);
            float t0, t1;
            const bool h0 = slab(r, a.x, a.y, a.z, a.w, c.x, c.y, tmin, best32, t0);
            const bool h1 = slab(r, b.x, b.y, b.z, b.w, c.z, c.w, tmin, best32, t1);
            const int c0 = (int)f2u_bits(m.x), c1 = (int)f2u_bits(m.y);
            if (h0 && h1) {
                const bool swap = t1 < t0;
                node = swap ? c1 : c0;
                stack[sp++] = swap ? c0 : c1;
            } else if (h0) {
                node = c0;
            } else if (h1) {
                node = c1;
            } else {
                node = stack[--sp];
            }
        } else {
            const int slot = ~node;
            if (STATS) stats->tris++;
            const TriRecord* tp = tris + slot;
            const float4 ta = B2PT_LDG4(&tp->a);
            const float4 tb = B2PT_LDG4(&tp->b);
            const float4 tc = B2PT_LDG4(&tp->c);
            const float3 v0 = v3(ta.x, ta.y, ta.z), v1 = v3(ta.w, tb.x, tb.y), v2 = v3(tb.z, tb.w, tc.x);
            if (tri_filter(r.o, r.d, r.d_l1, v0, v1, v2, tmin, best32)) {
                if (STATS) stats->exact++;
                double t, u, v;
                bool bf;
                if (tri_exact(r.o, r.d, v0, v1, v2, t, u, v, bf) && t > tmin_d && t < tmax_d) {
                    if (ANY) {
                        if (accept((uint32_t)slot, f2u_bits(tc.z), (float)u, (float)v)) {
                            hit.tri = (uint32_t)slot;
                            return true;
                        }
                        node = stack[--sp];
                        continue;
                    }
                    const uint32_t prim = f2u_bits(tc.y);
                    if ((t < best_t || (t == best_t && prim < best_prim)) &&
                        accept((uint32_t)slot, f2u_bits(tc.z), (float)u, (float)v)) {
                        best_t = t;
                        best_prim = prim;
                        // round up so the fp32 culling bound never cuts below best_t
                        float tf = (float)t;
                        if ((double)tf < t) tf = nextafterf(tf, INFINITY);
                        best32 = tf;
                        hit.tri = (uint32_t)slot | (bf ? kHitBackface : 0u);
                        hit.t = (float)t;
                        hit.u = (float)u;
                        hit.v = (float)v;
                    }
                }
            }
            node = stack[--sp];
        }
    }
    return hit.tri != kHitMiss;
}

}  // namespace b2pt
