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
#include <cmath>
#include <cstdint>

namespace {

struct Hit {
    bool hit;
    double t, u, v;
    bool backface;
};

inline Hit exact_triangle(const float* o, const float* d, const float* tri) {
    Hit h{false, 0, 0, 0, false};
    const double ox = o[0], oy = o[1], oz = o[2];
    const double dx = d[0], dy = d[1], dz = d[2];
    const double ax = tri[0], ay = tri[1], az = tri[2];
    const double e1x = (double)tri[3] - ax, e1y = (double)tri[4] - ay, e1z = (double)tri[5] - az;
    const double e2x = (double)tri[6] - ax, e2y = (double)tri[7] - ay, e2z = (double)tri[8] - az;
    const double px = dy * e2z - dz * e2y;
    const double py = dz * e2x - dx * e2z;
    const double pz = dx * e2y - dy * e2x;
    const double det = (e1x * px + e1y * py) + e1z * pz;
    if (det == 0.0) return h;
    const double inv = 1.0 / det;
    const double tx = ox - ax, ty = oy - ay, tz = oz - az;
    const double u = ((tx * px + ty * py) + tz * pz) * inv;
    if (!(u >= 0.0) || u > 1.0) return h;
    const double qx = ty * e1z - tz * e1y;
    const double qy = tz * e1x - tx * e1z;
    const double qz = tx * e1y - ty * e1x;
    const double v = ((dx * qx + dy * qy) + dz * qz) * inv;
    if (!(v >= 0.0) || u + v > 1.0) return h;
    const double t = ((e2x * qx + e2y * qy) + e2z * qz) * inv;
    h.hit = true;
    h.t = t;
    h.u = u;
    h.v = v;
    h.backface = det < 0.0;
    return h;
}

}  // namespace

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
