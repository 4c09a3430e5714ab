// optix.h — stand-in for the NVIDIA OptiX SDK header, TEST INFRASTRUCTURE ONLY.
// Lets the reference's device code (Renderer/OptiX/devicePrograms.cu and the
// headers it includes) compile as plain host C++ under g++.  Only the handful
// of types and enumerators that translation unit names are provided; see
// ../ref_harness.cpp for the semantics behind them.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <cmath>
using std::isinf;
using std::isnan;

typedef unsigned long long OptixTraversableHandle;
typedef unsigned int OptixVisibilityMask;
typedef unsigned long long CUdeviceptr_shim;
enum OptixRayFlags {
    OPTIX_RAY_FLAG_NONE = 0,
    OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1 << 0,
    OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT = 1 << 2,
};
typedef int OptixResult;
enum { OPTIX_SUCCESS = 0 };
