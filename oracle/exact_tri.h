// exact_tri.h — TEST INFRASTRUCTURE ONLY.  The fp64 Moeller-Trumbore decision that
// defines a hit (see the header of oracle_trace.cpp for the full contract).
// Every operation is one IEEE double operation in the written order; build with
// -ffp-contract=off.
#pragma once
#include <cmath>
#include <cstdint>

namespace oracle {


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


}  // namespace oracle
