// renderer.h — the object behind the opaque B2ptRenderer handle.
//
// Holds what the reference's OptixRenderer holds (OptixRenderer.h:10-61:
// context/stream, launch params, colour buffer, light buffer, per-mesh device
// arrays, the acceleration structure, textures) minus everything OptiX
// (module, pipeline, program groups, SBT), plus the wavefront work buffers.
#pragma once

#include <string>
#include <vector>

#include "wavefront.cuh"

struct B2ptRenderer {
    B2ptConfig cfg;
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string last_error;

    b2pt::SceneBuffers scene;

    // launch parameters (LaunchParams.h:9-29)
    int width = 0, height = 0;
    uint32_t frame_id = 0;
    bool have_camera = false;
    float cam_pos[3] = {0, 0, 0};
    float cam_inv_view[16] = {0};
    float cam_inv_proj[16] = {0};
    std::vector<B2ptPointLight> lights;
    b2pt::DeviceBuffer<B2ptPointLight> d_lights;

    // accumulation (replaces the GL framebuffer of OptixView.cpp:212-255)
    b2pt::DeviceBuffer<float> d_accum;     // W*H*3 fp32 sums
    uint32_t spp_accumulated = 0;

    // wavefront work buffers, sized for `wave_capacity` paths
    uint32_t wave_capacity = 0;
    b2pt::DeviceBuffer<float> d_words;         // backing store of all SoA float/u32 arrays
    b2pt::DeviceBuffer<b2pt::HitRecord> d_hits;
    b2pt::DeviceBuffer<uint32_t> d_queues;     // 4 material queues
    b2pt::DeviceBuffer<uint32_t> d_queues2;    // ... ordered by light visibility (K7p)
    b2pt::DeviceBuffer<uint32_t> d_counts;     // per-bounce counters
    b2pt::DeviceBuffer<float> d_wave_out;      // per-path radiance of the current wave (3 planes)
    b2pt::DeviceBuffer<float> d_layered;       // scratch of the layered walk engine (shade_layered.cu), allocated on first use
    b2pt::DeviceBuffer<b2pt::TraceCounters> d_counters;
    b2pt::DeviceBuffer<float> d_frame;         // W*H*3 single-frame output (Render)
    b2pt::PathState st[2];
    b2pt::SurfaceState surf;
    b2pt::ShadowRays shadow;

    // statistics
    uint64_t samples = 0;
    uint64_t kernel_launches = 0;
    uint64_t hook_radiance_rays = 0, hook_shadow_rays = 0;   // rays traced by the parity hooks
    double device_ms_render = 0.0;

    // per-kernel-class timing (b2pt_set_kernel_timing)
    bool kernel_timing = false;
    std::vector<cudaEvent_t> timing_events;   // pool, reused by every render call
    std::vector<int> timing_tags;             // class of the kernel that ended at event i+1
    size_t timing_used = 0;
    double ms_kernel[b2pt::kClsCount] = {0, 0, 0, 0, 0, 0};
    uint64_t launches_kernel_cls[b2pt::kClsCount] = {0, 0, 0, 0, 0, 0};

    // one process per device (multi.cu, b2pt_dist_*): the NCCL communicator this renderer reduces over
    void* dist_comm = nullptr;   // ncclComm_t
    int dist_rank = 0, dist_world = 1;
    b2pt::DeviceBuffer<unsigned long long> d_dist_count;   // [0] this rank's samples, [1] their sum on the root

    void set_error(const std::string& m) { last_error = m; }
};
