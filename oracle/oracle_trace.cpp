// oracle_trace.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// CPU statement of the ray/triangle contract of the path-tracing core.
//
// Parity status: UNPINNED for this file.  The reference delegates ray-triangle
// intersection to NVIDIA OptiX 7.x (closed source; SDK pinned at
// "OptiX SDK 7.3.0" in OptixPathtracer/OptixPathtracer.vcxproj:168; call sites
// optixTrace at Renderer/OptiX/devicePrograms.cu:224,248 and
// optixAccelBuild/Compact at OptixRenderer.cpp:416,440).  Its arithmetic is not
// published and the reference holds no golden vectors at that boundary, so the
// contract is DEFINED here (BASELINE.json north_star: "closest-hit primitive IDs
// and t-values bit-exact against a host double-precision brute-force oracle"):
//
//   * triangles are the fp32 world-space vertices (model matrix applied in
//     fp32, glm association, no FMA — scenes.scene_triangle_soup);
//   * a ray hits a triangle iff double-precision Moeller-Trumbore accepts it:
//     det != 0, 0 <= u <= 1, v >= 0, u + v <= 1, tmin < t < tmax  (no culling,
//     like OptiX with OPTIX_RAY_FLAG_NONE);
//   * closest hit = minimum double t; ties -> lower primitive index;
//   * back face iff det < 0 (front = counter-clockwise seen from the origin,
//     the convention of optixIsTriangleBackFaceHit, devicePrograms.cu:349);
//   * reported t,u,v are the fp32 roundings of the double values.
//
// Build: g++ -O2 -ffp-contract=off -fopenmp (see Makefile): every operation
// below is one IEEE double operation in the written order.
#include "exact_tri.h"
using oracle::Hit;
using oracle::exact_triangle;

extern "C" {

// Closest hit by exhaustive search.  tris: nt*9 floats (v0,v1,v2).
// Outputs per ray: prim (-1 = miss), t,u,v (fp32 roundings), backface (0/1).
void oracle_closest_bruteforce(const float* tris, int64_t nt, const float* org, const float* dir,
                               int64_t n, float tmin, float tmax, int32_t* prim, float* t, float* u,
                               float* v, uint8_t* backface) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; i++) {
        double best = (double)tmax;
        int64_t bp = -1;
        Hit bh{false, 0, 0, 0, false};
        for (int64_t k = 0; k < nt; k++) {
            const Hit h = exact_triangle(org + 3 * i, dir + 3 * i, tris + 9 * k);
            if (!h.hit || !(h.t > (double)tmin) || !(h.t < (double)tmax)) continue;
            if (h.t < best || (h.t == best && bp >= 0 && k < bp)) {
                best = h.t;
                bp = k;
                bh = h;
            }
        }
        prim[i] = (int32_t)bp;
        t[i] = bp >= 0 ? (float)bh.t : 0.0f;
        u[i] = bp >= 0 ? (float)bh.u : 0.0f;
        v[i] = bp >= 0 ? (float)bh.v : 0.0f;
        if (backface) backface[i] = bp >= 0 && bh.backface;
    }
}

// Any hit in (tmin, tmax[i]) by exhaustive search.
void oracle_any_bruteforce(const float* tris, int64_t nt, const float* org, const float* dir,
                           int64_t n, float tmin, const float* tmax, uint8_t* occluded) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; i++) {
        uint8_t occ = 0;
        for (int64_t k = 0; k < nt && !occ; k++) {
            const Hit h = exact_triangle(org + 3 * i, dir + 3 * i, tris + 9 * k);
            if (h.hit && h.t > (double)tmin && h.t < (double)tmax[i]) occ = 1;
        }
        occluded[i] = occ;
    }
}

// Second-closest distance bookkeeping for the documented tie rule: for each
// ray, how many triangles are hit at exactly the winning t (>= 1 on a hit).
void oracle_tie_count(const float* tris, int64_t nt, const float* org, const float* dir, int64_t n,
                      float tmin, float tmax, const int32_t* prim, int32_t* ties) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; i++) {
        ties[i] = 0;
        if (prim[i] < 0) continue;
        const Hit w = exact_triangle(org + 3 * i, dir + 3 * i, tris + 9 * (int64_t)prim[i]);
        for (int64_t k = 0; k < nt; k++) {
            const Hit h = exact_triangle(org + 3 * i, dir + 3 * i, tris + 9 * k);
            if (h.hit && h.t > (double)tmin && h.t < (double)tmax && h.t == w.t) ties[i]++;
        }
    }
}

}  // extern "C"

// ---- the host BVH (host_bvh.h), exposed so that tests can check it against
// the exhaustive search above -------------------------------------------------
#include "host_bvh.h"
extern "C" {
void oracle_closest_bvh(const float* tris, int64_t nt, const float* org, const float* dir, int64_t n,
                        float tmin, float tmax, int32_t* prim, float* t, float* u, float* v,
                        uint8_t* backface) {
    oracle::HostBvh bvh;
    bvh.build(tris, nt);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        const oracle::ClosestHit h = bvh.closest(org + 3 * i, dir + 3 * i, tmin, tmax);
        prim[i] = (int32_t)h.prim;
        t[i] = h.prim >= 0 ? (float)h.t : 0.0f;
        u[i] = h.prim >= 0 ? (float)h.u : 0.0f;
        v[i] = h.prim >= 0 ? (float)h.v : 0.0f;
        if (backface) backface[i] = h.prim >= 0 && h.backface;
    }
}
void oracle_any_bvh(const float* tris, int64_t nt, const float* org, const float* dir, int64_t n,
                    float tmin, const float* tmax, uint8_t* occluded) {
    oracle::HostBvh bvh;
    bvh.build(tris, nt);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) occluded[i] = bvh.any(org + 3 * i, dir + 3 * i, tmin, tmax[i]) ? 1 : 0;
}
}
