// host_demo.cpp — the INTEGRATION.md recipe, executed: a C++ host written against the reference's OWN headers
// and host sources (ModelLoading/ModelLoader, Renderer/Camera, GlmHelperMethods, Renderer/OptiX/LightsStruct.h,
// Renderer/Images/WriteImage — compiled in place from /root/reference by oracle/Makefile) with
// `#include "OptiX/OptixRenderer.h"` swapped for include/OptixRenderer.hpp and linked against libb2pt.so.
// It does what main.cpp:88-116 + OptixView::DrawOptix (OptixView.cpp:201-210) do, minus the window:
//
//   host_demo <folder/> <file.gltf> <W> <H> <bounces> <n_frames> <out.raw> <out.exr>
//
// out.raw = n_frames frames of W*H glm::vec3 as Render() filled them; out.exr = the last frame through the
// reference's WriteImage::WriteEXR.  TEST INFRASTRUCTURE (tests/test_host_code.py runs it on the GPU box).
#include <math.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
using std::isnan;

#include "ModelLoading/ModelLoader.h"
#include "Renderer/Camera.h"
#include "GlmHelperMethods.h"
#include "Renderer/OptiX/LightsStruct.h"
#include "Renderer/Images/WriteImage.h"

#include "OptixRenderer.hpp"
using OptixRenderer = b2pt_shim::OptixRendererT<Model>;   // the swap INTEGRATION.md describes

int main(int argc, char** argv) {
    if (argc < 9) {
        fprintf(stderr, "usage: host_demo folder/ file.gltf W H bounces frames out.raw out.exr\n");
        return 2;
    }
    const glm::ivec2 viewSize{atoi(argv[3]), atoi(argv[4])};
    const int maxBounces = atoi(argv[5]), frames = atoi(argv[6]);

    // main.cpp:13-18, 88-104 (Scene1's camera and its four lights)
    std::unique_ptr<Camera> camera(new Camera());
    camera->SetBlenderPosition(glm::vec3(3.85382f, 0.0f, 1.0f));
    camera->SetBlenderRotation(glm::vec3(90.0f, 0.0f, 90.0f));
    std::vector<PointLight> pointLights;
    const glm::vec3 lightColor(1);
    pointLights.push_back(PointLight{GlmHelper::BlenderToEnginePosition(1.33906f, -0.7f, 0.299367f), lightColor});
    pointLights.push_back(PointLight{GlmHelper::BlenderToEnginePosition(1.33906f, 0.7f, 0.299367f), lightColor});
    pointLights.push_back(PointLight{GlmHelper::BlenderToEnginePosition(1.33906f, 0.7f, 1.69937f), lightColor});
    pointLights.push_back(PointLight{GlmHelper::BlenderToEnginePosition(1.33906f, -0.7f, 1.69937f), lightColor});

    std::unique_ptr<Model> model = ModelLoader::LoadModel(argv[1], argv[2]);

    B2ptConfig cfg;
    b2pt_default_config(&cfg);
    cfg.rng_mode = B2PT_RNG_LCG_HOST;   // the reference's random sequence, so the frames can be compared bit for bit
    try {
        OptixRenderer renderer("source/Renderer/OptiX/devicePrograms.cu.ptx", model.get(), &cfg);   // OptixView.cpp:93
        glm::ivec2 size = viewSize;
        renderer.Resize(size);                       // OptixView.cpp:189
        renderer.SetLights(&pointLights);            // main.cpp:111
        renderer.SetMaxBounces(maxBounces);          // main.cpp:112
        std::vector<glm::vec3> pixels((size_t)viewSize.x * viewSize.y);
        FILE* raw = fopen(argv[7], "wb");
        if (!raw) return 3;
        for (int k = 0; k < frames; k++) {
            renderer.SetCamera(camera.get());        // OptixView.cpp:136
            renderer.Render(pixels.data());          // OptixView.cpp:207
            fwrite(pixels.data(), sizeof(glm::vec3), pixels.size(), raw);
        }
        fclose(raw);
        WriteImage::WriteEXR(pixels, viewSize, argv[8]);   // the reference's own writer on the last frame
    } catch (const std::exception& e) {
        fprintf(stderr, "host_demo: %s\n", e.what());
        return 1;
    }
    printf("host_demo ok: %d meshes, %d textures, %d frames of %dx%d\n", (int)model->meshes.size(),
           (int)model->textures.size(), frames, viewSize.x, viewSize.y);
    return 0;
}
