// vecmath.cuh — native float3 helpers for the path-tracing core.
//
// Replaces the reference's glm shims (Renderer/OptiX/glmCUDA.h:8-128,
// OptixHelpers.h:6-32).  Everything here is `B2PT_HD` so that the per-thread
// logic of every kernel can also be instantiated by the host-side test harness
// (tests/hostsim) — that harness is test infrastructure only; the product path
// always runs these functions inside CUDA kernels.
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define B2PT_HD __host__ __device__ __forceinline__
#define B2PT_D __device__ __forceinline__
#define B2PT_HD_CALL __host__ __device__ __noinline__ inline   // a real call (code size over call overhead)
#else
#define B2PT_HD_CALL inline
#define B2PT_HD inline
#define B2PT_D inline
#endif

namespace b2pt {

constexpr float kPi = 3.14159265359f;          // literal used throughout PBRT/*.h
constexpr float kInvPi = 0.31830988618379067154f;

B2PT_HD float3 v3(float x, float y, float z) { return make_float3(x, y, z); }
B2PT_HD float3 v3(float s) { return make_float3(s, s, s); }
B2PT_HD float3 operator+(float3 a, float3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
B2PT_HD float3 operator-(float3 a, float3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
B2PT_HD float3 operator-(float3 a) { return v3(-a.x, -a.y, -a.z); }
B2PT_HD float3 operator*(float3 a, float3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
B2PT_HD float3 operator*(float3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
B2PT_HD float3 operator*(float s, float3 a) { return v3(a.x * s, a.y * s, a.z * s); }
B2PT_HD float3 operator/(float3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
B2PT_HD float3 operator/(float3 a, float3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
B2PT_HD float3& operator+=(float3& a, float3 b) { a = a + b; return a; }
B2PT_HD float3& operator*=(float3& a, float3 b) { a = a * b; return a; }
B2PT_HD float3& operator*=(float3& a, float s) { a = a * s; return a; }
B2PT_HD float3& operator/=(float3& a, float s) { a = a / s; return a; }

B2PT_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
B2PT_HD float3 cross(float3 a, float3 b) {
    return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
B2PT_HD float length(float3 a) { return sqrtf(dot(a, a)); }
// glm::normalize = v * inversesqrt(dot(v,v)), where glm's scalar inversesqrt is
// 1/sqrt (glm/detail/func_exponential.inl) — kept as a true division so results
// agree with the reference to the last bits.
B2PT_HD float3 normalize(float3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
B2PT_HD float sqr(float x) { return x * x; }
B2PT_HD float absdot(float3 a, float3 b) { return fabsf(dot(a, b)); }
B2PT_HD float length_sqr(float3 a) { return sqr(a.x) + sqr(a.y) + sqr(a.z); }
B2PT_HD float max3(float3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }
B2PT_HD bool is_zero(float3 a) { return a.x == 0.0f && a.y == 0.0f && a.z == 0.0f; }
B2PT_HD float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
B2PT_HD bool any_nonfinite(float3 a) {
    return !(isfinite(a.x) && isfinite(a.y) && isfinite(a.z));
}

// Column-major 4x4 times (x,y,z,w), in glm's association order
// (glm/detail/type_mat4x4.inl operator*(mat4, vec4): (m0*x + m1*y) + (m2*z + m3*w)).
// Written without FMA contraction so that the world-space vertices (the input
// of the hit contract) are plain IEEE fp32 and any host can reproduce them.
#if defined(__CUDA_ARCH__)
#define B2PT_FMUL(a, b) __fmul_rn((a), (b))
#define B2PT_FADD(a, b) __fadd_rn((a), (b))
#else
#define B2PT_FMUL(a, b) ((a) * (b))
#define B2PT_FADD(a, b) ((a) + (b))
#endif
B2PT_HD float mat4_row(const float* m, int r, float x, float y, float z, float w) {
    return B2PT_FADD(B2PT_FADD(B2PT_FMUL(m[r], x), B2PT_FMUL(m[4 + r], y)),
                     B2PT_FADD(B2PT_FMUL(m[8 + r], z), B2PT_FMUL(m[12 + r], w)));
}
B2PT_HD float4 mat4_mul(const float* m, float x, float y, float z, float w) {
    float4 r;
    r.x = mat4_row(m, 0, x, y, z, w);
    r.y = mat4_row(m, 1, x, y, z, w);
    r.z = mat4_row(m, 2, x, y, z, w);
    r.w = mat4_row(m, 3, x, y, z, w);
    return r;
}

}  // namespace b2pt
