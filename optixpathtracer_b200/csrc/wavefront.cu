// wavefront.cu — placeholder for slice 1 (replaced by the integrator).
#include "render.h"
namespace b2pt {
void render_frames(B2ptRenderer*, int, bool) { throw Error{B2PT_E_UNSUPPORTED, "integrator not built in this slice"}; }
void resolve_mean(B2ptRenderer*) {}
}  // namespace b2pt
extern "C" {
int b2pt_bsdf_f(int32_t, int32_t, int32_t, int64_t, const float*, const float*, const float*, const float*, const uint32_t*, float*, uint32_t*) { return B2PT_E_UNSUPPORTED; }
int b2pt_bsdf_sample(int32_t, int32_t, int32_t, int64_t, const float*, const float*, const float*, const uint32_t*, uint8_t*, float*, float*, float*, uint8_t*, uint32_t*) { return B2PT_E_UNSUPPORTED; }
int b2pt_bsdf_pdf(int32_t, int32_t, int64_t, const float*, const float*, const float*, int32_t, int32_t, float*) { return B2PT_E_UNSUPPORTED; }
}
