// OptixRenderer.hpp — header-only C++ host class with the reference's surface
// (Renderer/OptiX/OptixRenderer.h:63-76) over the C-ABI of libb2pt.so.
//
// A maintainer of the reference replaces `#include "OptiX/OptixRenderer.h"` in
// Renderer/OptixView.h by this header and links libb2pt.so; OptixView.cpp:93
// (`new OptixRenderer(ptxPath, model)`), :136 (SetCamera), :207 (Render),
// :189/:194 (Resize) and main.cpp:111-112 (SetLights, SetMaxBounces) compile
// unchanged.  The template parameters keep this header free of the reference's
// own types: ModelT needs `meshes` (vector of pointers with vertecies / normal /
// texCoord / index / albedo / metallic / roughness / albedoTex / normalTex /
// metalRoughTex / GetModelMatrix(), ModelLoading/Mesh.h:9-47) and `textures`
// (pixel, resolution; Texture.h:5-13); CameraT needs position, GetViewMatrix(),
// GetProjectionMatrix(aspect) (Renderer/Camera.h).  Errors throw
// std::runtime_error like the reference's CUDA_CHECK (optix7.h:26-36).
#pragma once

#include <stdexcept>
#include <string>
#include <vector>

#include "b2pt.h"

namespace b2pt_shim {

inline void check(B2ptRenderer* r, int rc) {
    if (rc != B2PT_OK) throw std::runtime_error(std::string("b2pt: ") + b2pt_last_error(r));
}

template <class ModelT>
class OptixRendererT {
  public:
    ModelT* model;   // public field of the reference class (OptixRenderer.h:64)

    // ptxPath is accepted and ignored: there is no PTX to JIT.
    OptixRendererT(const std::string& /*ptxPath*/, ModelT* m, const B2ptConfig* cfg = nullptr) : model(m) {
        std::vector<B2ptMesh> meshes(m->meshes.size());
        for (size_t i = 0; i < meshes.size(); i++) {
            auto& src = *m->meshes[i];
            B2ptMesh& d = meshes[i];
            d.positions = src.vertecies.empty() ? nullptr : &src.vertecies[0].x;
            d.normals = src.normal.empty() ? nullptr : &src.normal[0].x;
            d.texcoords = src.texCoord.empty() ? nullptr : &src.texCoord[0].x;
            d.indices = src.index.empty() ? nullptr : &src.index[0].x;
            d.n_vertices = (int32_t)src.vertecies.size();
            d.n_triangles = (int32_t)src.index.size();
            const auto mm = src.GetModelMatrix();               // glm::mat4, column-major
            for (int k = 0; k < 16; k++) d.model[k] = (&mm[0][0])[k];
            d.albedo[0] = src.albedo.x; d.albedo[1] = src.albedo.y; d.albedo[2] = src.albedo.z;
            d.metallic = src.metallic;
            d.roughness = src.roughness;
            d.tex_albedo = src.albedoTex; d.tex_normal = src.normalTex; d.tex_metal_rough = src.metalRoughTex;
        }
        std::vector<B2ptTexture> textures(m->textures.size());
        for (size_t i = 0; i < textures.size(); i++) {
            auto& t = *m->textures[i];
            textures[i].rgba = reinterpret_cast<const uint8_t*>(t.pixel);
            textures[i].width = t.resolution.x;
            textures[i].height = t.resolution.y;
        }
        B2ptSceneDesc desc{meshes.data(), (int32_t)meshes.size(), textures.data(), (int32_t)textures.size()};
        int rc = b2pt_create(&desc, cfg, &r_);
        if (rc != B2PT_OK) throw std::runtime_error(std::string("b2pt_create: ") + b2pt_last_error(nullptr));
    }
    ~OptixRendererT() { b2pt_destroy(r_); }
    OptixRendererT(const OptixRendererT&) = delete;
    OptixRendererT& operator=(const OptixRendererT&) = delete;

    template <class IVec2>
    void Resize(IVec2& newSize) { w_ = newSize.x; h_ = newSize.y; check(r_, b2pt_resize(r_, newSize.x, newSize.y)); }

    // h_pixels: W*H glm::vec3 (12 bytes each, tight) — exactly what the reference fills
    template <class Vec3>
    void Render(Vec3 h_pixels[]) { check(r_, b2pt_render(r_, reinterpret_cast<float*>(h_pixels))); }

    // OptixRenderer::SetCamera (OptixRenderer.cpp:662-668) inverts view and projection on the host
    template <class CameraT>
    void SetCamera(CameraT* camera) {
        const float aspect = h_ ? float(w_) / float(h_) : 1.0f;
        const auto iv = inverse(camera->GetViewMatrix());
        const auto ip = inverse(camera->GetProjectionMatrix(aspect));
        check(r_, b2pt_set_camera(r_, &camera->position.x, &iv[0][0], &ip[0][0]));
    }

    template <class PointLightT>
    void SetLights(std::vector<PointLightT>* lights) {
        static_assert(sizeof(PointLightT) == sizeof(B2ptPointLight), "PointLight must be position, color (24 bytes)");
        check(r_, b2pt_set_lights(r_, reinterpret_cast<const B2ptPointLight*>(lights->data()), (int32_t)lights->size()));
    }

    void SetMaxBounces(int maxBounces) { check(r_, b2pt_set_max_bounces(r_, maxBounces)); }

    // extensions (the reference edits source for these)
    void SetBsdfMode(int mode) { check(r_, b2pt_set_bsdf_mode(r_, mode)); }
    void SetRng(int mode, uint64_t seed) { check(r_, b2pt_set_rng(r_, mode, seed)); }
    template <class Vec3>
    void RenderAccumulate(int spp, Vec3 h_mean[]) { check(r_, b2pt_render_accumulate(r_, spp, reinterpret_cast<float*>(h_mean))); }
    void ClearAccumulation() { check(r_, b2pt_clear_accumulation(r_)); }
    // WriteImage::WriteEXR / WriteBMP (Renderer/Images/WriteImage.cpp:9-99) and the Final.frag display
    // transform, applied to the accumulated image without leaving the library
    void WriteEXR(const char* path) { check(r_, b2pt_write_exr(r_, path)); }
    void WriteBMP(const char* path, bool apply_oetf = false) { check(r_, b2pt_write_bmp(r_, path, apply_oetf ? 1 : 0)); }
    void WritePNG(const char* path, bool apply_oetf = true) { check(r_, b2pt_write_png(r_, path, apply_oetf ? 1 : 0)); }
    B2ptRenderer* handle() const { return r_; }

  private:
    B2ptRenderer* r_ = nullptr;
    int w_ = 0, h_ = 0;
};

}  // namespace b2pt_shim

// In the reference tree:   #include "../ModelLoading/Model.h"
//                          using OptixRenderer = b2pt_shim::OptixRendererT<Model>;
