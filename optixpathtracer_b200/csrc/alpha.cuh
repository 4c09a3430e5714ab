// alpha.cuh — the alpha cut-out any-hit predicate (A9).
//
// __anyhit__radiance / __anyhit__shadow (devicePrograms.cu:549-561) call
// AlphaCutout (:518-547): on meshes with an albedo texture the texel at the
// interpolated uv is sRGB-decoded (all four channels, alpha included — SURVEY
// Q13) and the intersection is ignored when alpha < 0.9.  Meshes without
// texcoords read uv = (0, 0).
#pragma once

#include "integrator.cuh"

namespace b2pt {

template <class TexFetch>
struct AlphaCutout {
    DeviceScene sc;
    TexFetch tex;
    // meshes without an albedo texture accept every hit: the wide kernels keep their fp32-proof path for them
    B2PT_HD bool textured(uint32_t mesh) const { return sc.materials[mesh].tex_albedo >= 0; }
    B2PT_HD bool operator()(uint32_t slot, uint32_t mesh, float u, float v) const {
        const int t = sc.materials[mesh].tex_albedo;
        if (t < 0) return true;
        float tx = 0.0f, ty = 0.0f;
        if (sc.uvs) {
            const UvRecord ur = sc.uvs[slot];
            const float w0 = 1.0f - u - v;
            tx = w0 * ur.t0.x + u * ur.t1.x + v * ur.t2.x;
            ty = w0 * ur.t0.y + u * ur.t1.y + v * ur.t2.y;
        }
        const float a = srgb8_to_linear_1(tex(t, tx, ty).w);
        return !((double)a < 0.9);   // `fromTexture.a < 0.9` compares in double (:544)
    }
};

}  // namespace b2pt
