// scene.cuh — flattened device scene (what kernels see) and the build entry.
//
// The reference keeps one set of device arrays per mesh and reaches them per
// hit through the SBT record (MeshSBTData.h:6-29, filled in
// OptixRenderer.cpp:520-554): record -> index[prim] -> 3 x vertex / normal /
// texcoord, each vertex re-multiplied by modelMatrix in the closest-hit program
// (devicePrograms.cu:77-81).  Here everything is pre-transformed once and laid
// out per triangle in leaf order, so a hit costs 3+3 coalescable 16-byte loads
// and no matrix work.
#pragma once

#include <vector>

#include "bvh.cuh"
#include "host_util.h"

namespace b2pt {

// per-mesh material constants: MeshSBTData.albedoColor/roughness/metallic +
// texture ids (Mesh.h:29-37) + the first three columns of modelMatrix, which
// GetNormal applies to the interpolated object-space normal
// (devicePrograms.cu:96-100).  80 bytes, five LDG.128; the table has one entry
// per mesh and stays cache-resident.
struct alignas(16) MaterialRecord {
    float albedo[3];
    float metallic;
    float roughness;
    int32_t tex_albedo;   // < 0: none
    int32_t tex_normal;
    int32_t tex_metal_rough;
    float mcol[12];       // modelMatrix columns 0..2, 4 rows each (column-major)
};
static_assert(sizeof(MaterialRecord) == 80, "material record must be 80 bytes");

struct MeshInfo {
    float model[16];  // column-major
    int32_t vtx_offset;
    int32_t tri_offset;
    int32_t has_normals;
    int32_t has_uv;
    int32_t cutout;   // the mesh has an albedo texture whose alpha can reject a hit (some texel below 255): TriRecord::c.w
};

// POD view handed to kernels by value.
struct DeviceScene {
    const BvhNode2* nodes;
    const BvhNode4* nodes4;         // the same tree, two levels per node (bvh.cuh)
    const BvhNode4P* nodes4p;       // ... and those in 96 bytes, child indices hidden in the planes' low bytes (B2PT_STEP=6)
    const TriRecord* tris;
    const NrmRecord* nrms;
    const UvRecord* uvs;            // nullptr when no mesh has texcoords
    const MaterialRecord* materials;
    const cudaTextureObject_t* textures;
    int32_t n_nodes;                // binary nodes
    int32_t n_wide_nodes;           // wide nodes the traversal can reach (the compacted arrays above)
    int32_t n_tris;
    int32_t n_meshes;
    int32_t n_textures;
    int32_t any_alpha_texture;      // some mesh has an albedo texture (any-hit cut-out, A9)
    float max_abs;                  // largest |coordinate| of the geometry (the boxes are padded by 2^-17 of it)
};

struct SceneBuffers {
    DeviceBuffer<BvhNode2> nodes;
    DeviceBuffer<BvhNode4> nodes4;
    DeviceBuffer<BvhNode4P> nodes4p;
    DeviceBuffer<TriRecord> tris;
    DeviceBuffer<NrmRecord> nrms;
    DeviceBuffer<UvRecord> uvs;
    DeviceBuffer<MaterialRecord> materials;
    DeviceBuffer<cudaTextureObject_t> textures;
    DeviceBuffer<uint32_t> tri_order;   // leaf slot -> global primitive id
    std::vector<cudaArray_t> texture_arrays;
    std::vector<cudaTextureObject_t> texture_objects;
    int32_t n_tris = 0, n_nodes = 0, n_meshes = 0, n_textures = 0, any_alpha_texture = 0;
    int32_t n_wide_nodes = 0;           // reachable wide nodes, numbered breadth-first (K3e)
    int32_t depth = 0, wide_depth = 0;  // levels of the binary / wide tree, measured by the build
    float build_ms = 0.0f;
    float3 bounds_lo = {0, 0, 0}, bounds_hi = {0, 0, 0};

    DeviceScene view() const;
    int64_t device_bytes() const;
    void free_textures();
};

// what the selected traversal kernels need from the build (trace_kernels.cu): wide-node formats they read
// (bit 0 BvhNode4, bit 2 BvhNode4P — the build keeps only those), and whether a scene goes
// through the wide kernels or the binary ones (decides which stack bound the build checks)
int trace_wide_formats();
bool trace_uses_wide(int any_alpha_texture);

// Upload + flatten + LBVH build (K1-K3) on `stream`.  Throws b2pt::Error.  leaf_size = B2ptConfig::bvh_leaf_size:
// triangles per leaf of the wide tree (0 = the default, 1..8; bvh_build.cu `leafify`).
void build_scene(const B2ptSceneDesc& desc, SceneBuffers& out, cudaStream_t stream, int leaf_size = 0);

}  // namespace b2pt
