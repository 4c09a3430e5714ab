// ref_host.cpp — TEST INFRASTRUCTURE.  The reference's own HOST code either side of the hot path,
// compiled in place from /root/reference (nothing is copied) into oracle/_ref/libref_host.so:
//
//   N1  ModelLoading/ModelLoader.cpp:10-244 + Mesh.cpp (tinygltf + stb_image, vendored under 3rdParty/)
//       -> what `Model* model` holds after `ModelLoader::LoadModel(folder, file)`; pins
//       optixpathtracer_b200/gltf.py::load_gltf(compat=True)
//   N3  Renderer/Images/WriteImage.cpp:8-99 (WriteBMP / WriteEXR through stb_image_write / tinyexr)
//       -> the bytes of the files the reference writes; pins b2pt_write_bmp / b2pt_write_exr
//   --  Renderer/Camera.cpp + GlmHelperMethods.cpp -> view / projection of the Blender-style camera set-up of
//       main.cpp:83-84; pins optixpathtracer_b200/camera.py
//
// Only tests/ and tools/make_golden.py load this library.  The two WriteTexture* functions of WriteImage.cpp need
// an OpenGL texture; their GLTexture2D methods are stubbed to abort (never called).
#include <math.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
using std::isnan;

#include "ModelLoading/ModelLoader.cpp"
// ModelLoader.cpp emits the stb implementations (:4-7); the headers are pulled in again below (GLTexture2D.h,
// WriteImage.cpp) and their implementation sections carry no include guard
#undef STB_IMAGE_IMPLEMENTATION
#undef STB_IMAGE_WRITE_IMPLEMENTATION
#undef TINYGLTF_IMPLEMENTATION
#include "ModelLoading/Mesh.cpp"
#include "Renderer/Images/WriteImage.cpp"
#include "Renderer/Camera.cpp"
#include "GlmHelperMethods.cpp"

// GL-only members referenced by WriteTextureToBMP / WriteTextureToEXR (WriteImage.cpp:101-119)
void GLTexture2D::DownloadTexture(std::vector<glm::vec3>&) { abort(); }

extern "C" {

// ---- N1 ----------------------------------------------------------------------------------------
void* refhost_load_model(const char* folder, const char* file) {
    std::unique_ptr<Model> m = ModelLoader::LoadModel(folder, file);
    return m.release();
}
void refhost_model_free(void* h) {
    Model* m = static_cast<Model*>(h);
    if (!m) return;
    for (auto& t : m->textures) {   // stbi memory is malloc'ed; Texture::~Texture would delete[] it
        free(t->pixel);
        t->pixel = nullptr;
    }
    delete m;
}
void refhost_model_counts(void* h, int* n_meshes, int* n_textures) {
    Model* m = static_cast<Model*>(h);
    *n_meshes = (int)m->meshes.size();
    *n_textures = (int)m->textures.size();
}
// model_cm = GetModelMatrix() column-major (what OptixRenderer uploads transposed, OptixRenderer.cpp:352-353)
void refhost_mesh_info(void* h, int i, int* n_vert, int* n_tri, int* n_uv, float* model_cm, float* albedo,
                       float* metallic, float* roughness, int* tex) {
    Mesh& me = *static_cast<Model*>(h)->meshes[i];
    *n_vert = (int)me.vertecies.size();
    *n_tri = (int)me.index.size();
    *n_uv = (int)me.texCoord.size();
    const glm::mat4 mm = me.GetModelMatrix();
    memcpy(model_cm, &mm[0][0], 64);
    albedo[0] = me.albedo.x; albedo[1] = me.albedo.y; albedo[2] = me.albedo.z;
    *metallic = me.metallic;
    *roughness = me.roughness;
    tex[0] = me.albedoTex; tex[1] = me.normalTex; tex[2] = me.metalRoughTex;
}
void refhost_mesh_arrays(void* h, int i, float* pos, float* nrm, float* uv, int32_t* idx) {
    Mesh& me = *static_cast<Model*>(h)->meshes[i];
    if (pos) memcpy(pos, me.vertecies.data(), me.vertecies.size() * 12);
    if (nrm) memcpy(nrm, me.normal.data(), me.normal.size() * 12);
    if (uv) memcpy(uv, me.texCoord.data(), me.texCoord.size() * 8);
    if (idx) memcpy(idx, me.index.data(), me.index.size() * 12);
}
void refhost_texture_info(void* h, int i, int* w, int* hgt) {
    Texture& t = *static_cast<Model*>(h)->textures[i];
    *w = t.resolution.x;
    *hgt = t.resolution.y;
}
void refhost_texture_pixels(void* h, int i, uint8_t* rgba) {
    Texture& t = *static_cast<Model*>(h)->textures[i];
    memcpy(rgba, t.pixel, (size_t)t.resolution.x * t.resolution.y * 4);
}

// ---- N3 ----------------------------------------------------------------------------------------
// rgb: W*H*3 floats, row 0 = bottom (what OptixRenderer::Render fills, what OptixView hands to WriteImage)
void refhost_write_exr(const float* rgb, int w, int h, const char* path) {
    std::vector<glm::vec3> px((size_t)w * h);
    memcpy(px.data(), rgb, px.size() * 12);
    WriteImage::WriteEXR(px, glm::ivec2(w, h), path);
}
void refhost_write_bmp(const float* rgb, int w, int h, const char* path) {
    std::vector<glm::vec3> px((size_t)w * h);
    memcpy(px.data(), rgb, px.size() * 12);
    WriteImage::WriteBMP(px, glm::ivec2(w, h), path);
}

// ---- camera ------------------------------------------------------------------------------------
// main.cpp:83-84: SetBlenderPosition / SetBlenderRotation, default FOV 40 (Camera.cpp:7); aspect from the frame
// size (OptixRenderer.cpp:663).  Outputs: position, view, projection (glm column-major).
void refhost_camera(const float* blender_pos, const float* blender_rot, float fov_deg, float aspect, float* pos_out,
                    float* view_cm, float* proj_cm) {
    Camera c;
    c.SetHorizontalFOV(fov_deg);
    c.SetBlenderPosition(glm::vec3(blender_pos[0], blender_pos[1], blender_pos[2]));
    c.SetBlenderRotation(glm::vec3(blender_rot[0], blender_rot[1], blender_rot[2]));
    pos_out[0] = c.position.x; pos_out[1] = c.position.y; pos_out[2] = c.position.z;
    const glm::mat4 v = c.GetViewMatrix(), p = c.GetProjectionMatrix(aspect);
    memcpy(view_cm, &v[0][0], 64);
    memcpy(proj_cm, &p[0][0], 64);
}
// engine-space set-up (position / rotation given directly)
void refhost_camera_engine(const float* pos, const float* rot, float fov_deg, float aspect, float* view_cm, float* proj_cm) {
    Camera c;
    c.SetHorizontalFOV(fov_deg);
    c.position = glm::vec3(pos[0], pos[1], pos[2]);
    c.rotation = glm::vec3(rot[0], rot[1], rot[2]);
    const glm::mat4 v = c.GetViewMatrix(), p = c.GetProjectionMatrix(aspect);
    memcpy(view_cm, &v[0][0], 64);
    memcpy(proj_cm, &p[0][0], 64);
}

}  // extern "C"
