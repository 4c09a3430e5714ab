// optix_device.h — host stand-ins for the OptiX device intrinsics used by
// Renderer/OptiX/devicePrograms.cu (13 distinct _optix_* symbols in the
// committed PTX).  TEST INFRASTRUCTURE ONLY.
//
// State that OptiX keeps per ray/thread lives in a thread_local record that
// ref_harness.cpp fills before it calls the reference's program functions.
#pragma once
#include "optix.h"
#include <cuda_runtime.h>

struct ShimThreadState {
    uint3 launch_index;
    // current hit (valid inside closest-hit / any-hit)
    float2 bary;
    unsigned int prim;
    bool backface;
    float3 ray_dir;
    void* sbt_data;
    // payload registers of the current trace
    uint32_t p0, p1;
    bool ignored;
};
extern thread_local ShimThreadState g_shim;

static inline uint3 optixGetLaunchIndex() { return g_shim.launch_index; }
static inline float2 optixGetTriangleBarycentrics() { return g_shim.bary; }
static inline unsigned int optixGetPrimitiveIndex() { return g_shim.prim; }
static inline bool optixIsTriangleBackFaceHit() { return g_shim.backface; }
static inline float3 optixGetWorldRayDirection() { return g_shim.ray_dir; }
static inline unsigned long long optixGetSbtDataPointer() { return (unsigned long long)g_shim.sbt_data; }
static inline uint32_t optixGetPayload_0() { return g_shim.p0; }
static inline uint32_t optixGetPayload_1() { return g_shim.p1; }
static inline void optixIgnoreIntersection() { g_shim.ignored = true; }

// defined in ref_harness.cpp after the reference's programs are visible
void optixTrace(OptixTraversableHandle handle, float3 origin, float3 direction, float tmin, float tmax,
                float rayTime, OptixVisibilityMask mask, unsigned int rayFlags, unsigned int sbtOffset,
                unsigned int sbtStride, unsigned int missSbtIndex, uint32_t& p0, uint32_t& p1);

// tex2D<float4>: bilinear, wrap, normalised coordinates, normalised-float reads
// of an RGBA8 image (the texture descriptor of OptixRenderer.cpp:592-603).
struct ShimTexture {
    const unsigned char* rgba;
    int width, height;
};
float4 shim_tex2d(unsigned long long obj, float x, float y);
template <typename T>
static inline T tex2D(unsigned long long obj, float x, float y) {
    return shim_tex2d(obj, x, y);
}
