/*
 * b2pt.h — C-ABI of the B200-native path-tracing core (libb2pt.so).
 *
 * This is the drop-in boundary for the hot path of Damo12320/OptixPathtracer:
 * the reference's only seam into its device programs is the C++ class
 * `OptixRenderer` (OptixPathtracer/source/Renderer/OptiX/OptixRenderer.h:8-110).
 * Every entry point below cites the reference member it replaces.  All
 * pointers are plain host pointers unless the name says `_device`; matrices are
 * column-major float[16] exactly like `glm::mat4` (LaunchParams.h:17-21).
 *
 * Conventions
 *   - every call returns B2PT_OK (0) or a negative B2PT_E_* code; nothing
 *     throws across the ABI and nothing calls exit() (the reference throws from
 *     CUDA_CHECK and exits from OPTIX_CHECK, 3rdParty/OptixSample/optix7.h:26-62);
 *   - one handle = one host thread at a time;
 *   - image layout: W*H*3 packed floats, pixel (x,y) at 3*(y*W+x), row 0 is
 *     the bottom row (launch index y, devicePrograms.cu:704-705).
 *   - there is NO CPU fallback: without a CUDA device every compute call
 *     returns B2PT_E_CUDA.
 */
#ifndef B2PT_H
#define B2PT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2PT_ABI_VERSION 4

/* ---- status codes ------------------------------------------------------ */
enum {
    B2PT_OK = 0,
    B2PT_E_INVALID = -1,   /* bad argument (null pointer, negative size, bad enum) */
    B2PT_E_CUDA = -2,      /* CUDA runtime error; see b2pt_last_error */
    B2PT_E_NOMEM = -3,     /* host or device allocation failed */
    B2PT_E_STATE = -4,     /* call order (e.g. render before resize/set_camera) */
    B2PT_E_UNSUPPORTED = -5
};

/* ---- BSDF selection ---------------------------------------------------- *
 * The reference picks the BSDF by editing source between renders
 * (devicePrograms.cu:303-341, commented alternatives at 304-307 / 324-327).
 * AUTO is the shipped behaviour: rnd < metallic ? Conductor : GlossyDiffuse
 * (devicePrograms.cu:400, 310-318, 329-336).                              */
enum {
    B2PT_BSDF_AUTO = 0,
    B2PT_BSDF_CONDUCTOR = 1,   /* PBRT/Conductor.h:97,122 */
    B2PT_BSDF_DIFFUSE = 2,     /* PBRT/LambertDiffuse.h:86,110 */
    B2PT_BSDF_DIELECTRIC = 3,  /* PBRT/Dielectric.h:96,146 */
    B2PT_BSDF_LAYERED = 4      /* PBRT/GlossyDiffuse.h:141,372 */
};

/* ---- random-number policy ---------------------------------------------- *
 * LCG_* reproduce the reference stream (random.h:34-69: tea<16> seed,
 * 24-bit LCG).  The two LCG flavours differ only where the reference's result
 * depends on the compiler: evaluation order of `glm::vec2(rnd(s), rnd(s))`
 * (random.h:79, LambertDiffuse.h:40) and the negative float->unsigned cast in
 * GlossyDiffuse.h:215-217,417.  DEVICE = what nvcc/PTX does (left-to-right,
 * saturating cast), HOST = what g++/x86-64 does (right-to-left, wrapping cast;
 * this is the flavour the host-compiled oracle produces).
 * PHILOX is the production generator: Philox4x32-10 keyed by
 * (seed, pixel, absolute sample index), counter = (bounce, draw block).     */
enum {
    B2PT_RNG_PHILOX = 0,
    B2PT_RNG_LCG_DEVICE = 1,
    B2PT_RNG_LCG_HOST = 2
};

/* ---- scene hand-off ---------------------------------------------------- *
 * Mirrors what OptixRenderer reads from `Model* model`
 * (ModelLoading/Mesh.h:9-47, Texture.h:5-13; upload in
 * OptixRenderer.cpp:328-340 and SBT fill :520-554).  The library copies
 * everything during b2pt_create; the caller may free afterwards.           */
typedef struct B2ptMesh {
    const float*   positions;   /* n_vertices * 3, object space              */
    const float*   normals;     /* n_vertices * 3, may be NULL (Ns = 0)      */
    const float*   texcoords;   /* n_vertices * 2, may be NULL               */
    const int32_t* indices;     /* n_triangles * 3 (glm::ivec3, Mesh.h:20)   */
    int32_t        n_vertices;
    int32_t        n_triangles;
    float          model[16];   /* column-major = Mesh::GetModelMatrix()     */
    float          albedo[3];
    float          metallic;
    float          roughness;
    int32_t        tex_albedo;      /* index into textures, <0 = none        */
    int32_t        tex_normal;
    int32_t        tex_metal_rough;
} B2ptMesh;

typedef struct B2ptTexture {
    const uint8_t* rgba;        /* width*height*4, row-major, as stbi_load   */
    int32_t        width;
    int32_t        height;
} B2ptTexture;

typedef struct B2ptSceneDesc {
    const B2ptMesh*    meshes;
    int32_t            n_meshes;
    const B2ptTexture* textures;
    int32_t            n_textures;
} B2ptSceneDesc;

/* PointLight, LightsStruct.h:6-10 (position, color; 24 bytes, tight). */
typedef struct B2ptPointLight {
    float position[3];
    float color[3];
} B2ptPointLight;

/* Everything the reference hard-codes (SURVEY §5 "Config / flags").
 * Zero-initialise, then override; b2pt_default_config fills the reference's
 * constants.                                                               */
typedef struct B2ptConfig {
    int32_t  device;          /* CUDA ordinal; reference: 0 (OptixRenderer.cpp:70) */
    int32_t  bsdf_mode;       /* B2PT_BSDF_*                                      */
    int32_t  rng_mode;        /* B2PT_RNG_*                                       */
    uint64_t seed;            /* Philox key; ignored by the LCG modes             */
    int32_t  max_bounces;     /* main.cpp:87 uses 2; b2pt default 8               */
    int32_t  wave_paths_log2; /* in-flight paths per wavefront (log2), 0 = auto   */
    int32_t  bvh_leaf_size;   /* triangles per leaf of the wide tree, 1..8; 0 = 2 */
    int32_t  collect_stats;   /* 1 = count nodes/triangles per ray (slower)       */
} B2ptConfig;

typedef struct B2ptStats {
    /* counters since the last b2pt_reset_stats */
    uint64_t samples;            /* camera paths completed                        */
    uint64_t radiance_rays;      /* closest-hit rays traced                       */
    uint64_t shadow_rays;        /* any-hit rays traced                           */
    uint64_t nodes_visited;      /* only with collect_stats                       */
    uint64_t triangles_tested;   /* only with collect_stats                       */
    uint64_t exact_tests;        /* fp64 decisions taken (only with collect_stats)*/
    uint64_t nan_pixels;         /* NaN/Inf samples zeroed (WriteImage.cpp:52-55) */
    uint64_t kernel_launches;    /* kernels launched by render calls              */
    double   device_ms_render;   /* CUDA-event time of render calls               */
    double   device_ms_build;    /* CUDA-event time of the last BVH build         */
    /* static facts */
    int32_t  n_triangles;
    int32_t  n_nodes;
    int32_t  node_bytes;
    int32_t  triangle_bytes;
    int64_t  scene_device_bytes;
    /* shadow-ray share of the traversal counters (only with collect_stats);
     * nodes_visited / triangles_tested / exact_tests above are the closest-hit
     * share */
    uint64_t nodes_visited_shadow;
    uint64_t triangles_tested_shadow;
    uint64_t exact_tests_shadow;
    /* per-kernel-class device time of render calls, CUDA events on the
     * renderer's stream (only while b2pt_set_kernel_timing is on):
     * [0] K5 closest-hit traversal, [1] K8 any-hit traversal, [2] K6 surface,
     * [3] K7 shade (all BSDFs), [4] K4 raygen + K9 accumulate + bookkeeping   */
    double   device_ms_kernel[5];
    uint64_t launches_kernel[5];
    /* shadow rays K8 really traced.  shadow_rays above is the reference's count (one per
     * shaded hit while lights exist); the difference are rays whose BSDF value towards the
     * light is exactly zero (mirror / smooth glass, light behind an opaque surface) and which
     * K6 therefore never emits — same image, fewer rays.             */
    uint64_t shadow_rays_traced;
    /* K5 with collect_stats: lane slots issued by the node loop and the triangle loop of the
     * warp-persistent kernel (32 per warp trip).  nodes_visited / warp_slots_node and
     * triangles_tested / warp_slots_tri are the SIMD utilisation of the two loops.          */
    uint64_t warp_slots_node;
    uint64_t warp_slots_tri;
    /* the walk engine of the layered BSDF (production generator; csrc/shade_layered.cu): device time and
     * launches of the engine kernel alone (both are also inside device_ms_kernel[3] / launches_kernel[3]),
     * hits whose f was evaluated (five walk tasks each) and hits that ran a sample walk (one task each)   */
    double   device_ms_engine;
    uint64_t launches_engine;
    uint64_t walk_f_hits;
    uint64_t walk_s_hits;
    /* acceleration structure as built: wide nodes the traversal can reach (the compacted array K5/K8 read),
     * levels of the binary and of the wide tree (the traversal stacks hold 64 and 3 x 64 entries)          */
    int32_t  n_wide_nodes;
    int32_t  bvh_depth;
    int32_t  bvh_wide_depth;
    int32_t  reserved0;
} B2ptStats;

typedef struct B2ptRenderer B2ptRenderer;

/* ---- life cycle -------------------------------------------------------- */
int  b2pt_abi_version(void);
void b2pt_default_config(B2ptConfig* cfg);

/* OptixRenderer::OptixRenderer(ptxPath, model) (OptixRenderer.cpp:9-36):
 * upload scene, build the acceleration structure (BuildAccel :306-456 ->
 * GPU LBVH here), textures (:562-612), material table (BuildSBT :486-560). */
int  b2pt_create(const B2ptSceneDesc* scene, const B2ptConfig* cfg, B2ptRenderer** out);
/* The reference has no destructor; this frees every device buffer.         */
void b2pt_destroy(B2ptRenderer* r);
/* Thread-safe-per-handle message of the last failing call (r may be NULL
 * to read the message of a failed b2pt_create).                            */
const char* b2pt_last_error(const B2ptRenderer* r);

/* ---- renderer surface (same names/meaning as the reference class) ------ */
/* OptixRenderer::Resize (OptixRenderer.cpp:649-660); 0-sized = silent no-op,
 * returns B2PT_OK like the reference (:651).  Clears accumulation.          */
int b2pt_resize(B2ptRenderer* r, int32_t width, int32_t height);
/* OptixRenderer::SetCamera (:662-668): the caller passes what the reference
 * stores in LaunchParams.camera — position, inverse(view), inverse(proj).   */
int b2pt_set_camera(B2ptRenderer* r, const float position[3],
                    const float inverse_view[16], const float inverse_projection[16]);
/* OptixRenderer::SetLights (:670-675); re-callable (the reference leaks).   */
int b2pt_set_lights(B2ptRenderer* r, const B2ptPointLight* lights, int32_t n_lights);
/* OptixRenderer::SetMaxBounces (:677). */
int b2pt_set_max_bounces(B2ptRenderer* r, int32_t max_bounces);
int b2pt_set_bsdf_mode(B2ptRenderer* r, int32_t bsdf_mode);
int b2pt_set_rng(B2ptRenderer* r, int32_t rng_mode, uint64_t seed);
/* launchParams.frame.id (LaunchParams.h:14): the per-frame RNG salt; also the
 * absolute index of the next sample.  Render calls pre-increment it exactly
 * like OptixRenderer::Render (:623).                                        */
int b2pt_set_frame_id(B2ptRenderer* r, uint32_t frame_id);
int b2pt_get_frame_id(const B2ptRenderer* r, uint32_t* frame_id);

/* OptixRenderer::Render(glm::vec3 h_pixels[]) (:617-647): frame.id++, ONE
 * sample per pixel, result overwrites h_rgb (W*H*3 floats), synchronous.
 * Silent no-op (B2PT_OK) before the first resize, like :621.               */
int b2pt_render(B2ptRenderer* r, float* h_rgb);

/* Extension replacing OptixView::AddNewFrameToBuffer + AddPathtracedFrame.frag
 * (OptixView.cpp:212-255, frag :18-24): `spp` further samples per pixel are
 * traced and summed on the device (frame.id advances by spp); h_rgb_mean
 * (may be NULL) receives sum/total_spp of everything accumulated since the
 * last b2pt_clear_accumulation, NaN/Inf samples zeroed and counted.         */
int b2pt_render_accumulate(B2ptRenderer* r, int32_t spp, float* h_rgb_mean);
/* Same, but nothing is downloaded: the un-normalised fp32 sum stays on the
 * device (multi-GPU: the caller reduces the sums, e.g. NCCL, then scales).  */
int b2pt_render_accumulate_device(B2ptRenderer* r, int32_t spp);
int b2pt_clear_accumulation(B2ptRenderer* r);
/* Device pointer to the W*H*3 fp32 sum buffer and the samples summed in it. */
int b2pt_get_accumulation_device(B2ptRenderer* r, float** d_rgb_sum, uint32_t* spp_accumulated);
/* Checkpoint/resume of (sum, spp).                              */
int b2pt_download_accumulation(B2ptRenderer* r, float* h_rgb_sum, uint32_t* spp_accumulated);
int b2pt_upload_accumulation(B2ptRenderer* r, const float* h_rgb_sum, uint32_t spp_accumulated);

/* ---- image output (replaces the GL display pass and WriteImage.cpp) ----
 * All of them act on the running mean sum/spp of the accumulation buffer.
 * b2pt_display_rgb8: Final.frag:12-36 (LinearToSRGB8 when apply_oetf != 0) then
 *   WriteImage.cpp:16-18 (clamp to [0,1], (uint8_t)(c * 255.0f)); h_rgb8 holds W*H*3
 *   bytes; top_row_first != 0 flips the rows (files), 0 keeps the API's row 0 = bottom.
 * b2pt_write_bmp / b2pt_write_png: the same 8-bit image as a file (WriteBMP :9-33;
 *   PNG: what the README's Images/PNGs hold).
 * b2pt_write_exr: WriteEXR :36-99 — FLOAT channels B, G, R, top row first, a pixel
 *   with a NaN component written as 0; scan-line OpenEXR without compression.    */
int b2pt_display_rgb8(B2ptRenderer* r, int32_t apply_oetf, int32_t top_row_first, uint8_t* h_rgb8);
int b2pt_write_bmp(B2ptRenderer* r, const char* path, int32_t apply_oetf);
int b2pt_write_png(B2ptRenderer* r, const char* path, int32_t apply_oetf);
int b2pt_write_exr(B2ptRenderer* r, const char* path);

/* Record a CUDA event after every kernel of the render calls and attribute the
 * elapsed device time per kernel class (B2ptStats.device_ms_kernel).         */
int b2pt_set_kernel_timing(B2ptRenderer* r, int32_t on);
int b2pt_get_stats(const B2ptRenderer* r, B2ptStats* out);
int b2pt_reset_stats(B2ptRenderer* r);

/* ---- multi-GPU -------------------------------------------- *
 * The reference renders on device 0 only (OptixRenderer.cpp:70).  Here the path
 * shards by SAMPLES: every device renders a disjoint range of the absolute sample
 * indices (the Philox key holds the absolute index, so the union is the 1-GPU
 * image up to fp32 re-association), scene + BVH are replicated, and the W*H*3
 * fp32 sums are combined once per call.
 *
 * (a) ONE process, N devices.  b2pt_create_multi builds one renderer per device
 *     (concurrently); the b2pt_multi_* setters broadcast; b2pt_multi_renderer gives
 *     the per-device handle for everything else (stats, per-device settings).
 *     b2pt_multi_render_accumulate splits `spp_total` over the devices, renders on
 *     one host thread per device, and sums the buffers with ONE kernel on
 *     devices[0] that reads its peers' memory over NVLink (device order, so the
 *     result is deterministic), scales by 1 / spp and downloads.               */
typedef struct B2ptMulti B2ptMulti;
int  b2pt_create_multi(const int32_t* devices, int32_t n_devices, const B2ptSceneDesc* scene,
                       const B2ptConfig* cfg /* .device ignored */, B2ptMulti** out);
void b2pt_destroy_multi(B2ptMulti* m);
const char* b2pt_multi_last_error(const B2ptMulti* m);
int32_t b2pt_multi_device_count(const B2ptMulti* m);
B2ptRenderer* b2pt_multi_renderer(B2ptMulti* m, int32_t i);
int b2pt_multi_resize(B2ptMulti* m, int32_t width, int32_t height);
int b2pt_multi_set_camera(B2ptMulti* m, const float position[3],
                          const float inverse_view[16], const float inverse_projection[16]);
int b2pt_multi_set_lights(B2ptMulti* m, const B2ptPointLight* lights, int32_t n_lights);
int b2pt_multi_set_max_bounces(B2ptMulti* m, int32_t max_bounces);
int b2pt_multi_set_bsdf_mode(B2ptMulti* m, int32_t bsdf_mode);
int b2pt_multi_set_rng(B2ptMulti* m, int32_t rng_mode, uint64_t seed);
int b2pt_multi_set_frame_id(B2ptMulti* m, uint32_t frame_id);
int b2pt_multi_clear_accumulation(B2ptMulti* m);
int b2pt_multi_render_accumulate(B2ptMulti* m, int32_t spp_total, float* h_rgb_mean);

/* (b) one process PER device (torchrun / MPI).  Rank 0 calls b2pt_dist_unique_id,
 *     the host program hands the 128 bytes to every rank (torch.distributed
 *     broadcast, MPI_Bcast, a file ...), every rank calls b2pt_dist_init on its own
 *     renderer (created with B2ptConfig.device = its GPU).  Then
 *     b2pt_dist_render_accumulate is a COLLECTIVE: this rank traces `spp` further
 *     samples starting at its own frame id (give the ranks disjoint ranges with
 *     b2pt_set_frame_id), the accumulation buffers are summed on rank 0 with one
 *     ncclReduce(sum, fp32, W*H*3) on the renderer's stream straight out of the
 *     buffer K9 wrote, and rank 0's h_rgb_mean (may be NULL; ignored elsewhere)
 *     receives sum / (samples accumulated over all ranks since their last clear).
 *     NCCL (libnccl.so.2) is dlopen'ed by the first b2pt_dist_* call; the library
 *     loads and renders without it.                                            */
#define B2PT_DIST_ID_BYTES 128
int b2pt_dist_unique_id(uint8_t id[B2PT_DIST_ID_BYTES]);
int b2pt_dist_init(B2ptRenderer* r, const uint8_t id[B2PT_DIST_ID_BYTES], int32_t rank, int32_t world);
int b2pt_dist_render_accumulate(B2ptRenderer* r, int32_t spp, float* h_rgb_mean);
int b2pt_dist_shutdown(B2ptRenderer* r);

/* ---- parity hooks (the same kernels the renderer launches) ------------- */
/* optixTrace(..., RADIANCE ...) (devicePrograms.cu:248): closest hit of n
 * rays over the scene.  org/dir are n*3 floats.  Outputs: prim = global
 * triangle index (meshes concatenated in order) or -1, mesh, t/u/v = fp32
 * roundings of the fp64 Moeller-Trumbore decision.          */
int b2pt_trace_closest(B2ptRenderer* r, int64_t n, const float* org, const float* dir,
                       float tmin, float tmax,
                       int32_t* prim, int32_t* mesh, float* t, float* u, float* v);
/* optixTrace(..., SHADOW, TERMINATE_ON_FIRST_HIT) (devicePrograms.cu:224):
 * occluded[i] = 1 iff some triangle has tmin < t < tmax[i].                 */
int b2pt_trace_any(B2ptRenderer* r, int64_t n, const float* org, const float* dir,
                   float tmin, const float* tmax, uint8_t* occluded);
/* b2pt_trace_any for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  No
 * counterpart in the reference.
 * Same arguments, same kernel, same answers; the ray planes, the ray count
 * and the answers live in one file-scope __device__ array (csrc/noalloc.cu,
 * 1.7 MB held by every GPU that loads the library until exit) instead of
 * per-call cudaMalloc'ed buffers, and the call makes no cudaMalloc / cudaFree
 * / allocating library call.  n > B2PT_NOALLOC_MAX_RAYS returns
 * B2PT_E_UNSUPPORTED (b2pt_last_error says so).  Every call on a device
 * shares the array: calls that overlap on one GPU race — serialise them.   */
#define B2PT_NOALLOC_MAX_RAYS 32768
int b2pt_trace_any_noalloc(B2ptRenderer* r, int64_t n, const float* org, const float* dir,
                           float tmin, const float* tmax, uint8_t* occluded);

/* BSDF device functions, one thread per row (PBRT/ headers, all in local
 * shading space, z = normal).  `seed` is the LCG state on entry (rng policy
 * = lcg_mode, B2PT_RNG_LCG_DEVICE or _HOST); seed_out (may be NULL) is the
 * state on return, so draw counts are checked too.
 * lcg_mode = B2PT_RNG_PHILOX is the production instantiation: the same
 * functions drawing from Philox keyed (seed[i], row); seed_out is zeroed.
 * With bsdf = B2PT_BSDF_LAYERED it runs the per-lane
 * logic of the production walk engine instead (csrc/fast_layered.cuh): seed
 * keys the Philox streams, seed_out is zeroed, and `sample` reports
 * f = weight / |wi.z| with pdf = 1 (the engine only forms f |cos| / pdf).
 *   f      : n*3 out
 *   sample : ok[n] (0/1), f n*3, wi n*3, pdf n, flags n
 *            (bit0 reflection, bit1 transmission, bit2 specular, bit3 glossy)
 *   pdf    : Dielectric::PDF / LambertDiffuse::PDF with (reflection,
 *            transmission) lobe switches                                    */
int b2pt_bsdf_f(int32_t device, int32_t bsdf, int32_t lcg_mode, int64_t n,
                const float* albedo, const float* roughness,
                const float* wo, const float* wi, const uint32_t* seed,
                float* f_out, uint32_t* seed_out);
int b2pt_bsdf_sample(int32_t device, int32_t bsdf, int32_t lcg_mode, int64_t n,
                     const float* albedo, const float* roughness,
                     const float* wo, const uint32_t* seed,
                     uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                     uint8_t* flags_out, uint32_t* seed_out);
int b2pt_bsdf_pdf(int32_t device, int32_t bsdf, int64_t n,
                  const float* roughness, const float* wo, const float* wi,
                  int32_t reflection, int32_t transmission, float* pdf_out);

/* Read back the built acceleration structure (tests: structural checks).   *
 * nodes: n_nodes * 64 raw bytes, the binary nodes the builder emits (the   *
 * 4-wide nodes the traversal walks, node_bytes = 128, are derived from      *
 * them); tri_order: n_triangles global triangle ids in leaf order.  Either  *
 * pointer may be NULL.                                                      */
int b2pt_download_bvh(B2ptRenderer* r, void* nodes, int32_t* tri_order);

#ifdef __cplusplus
}
#endif
#endif /* B2PT_H */
