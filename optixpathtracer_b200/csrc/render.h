// render.h — host-side entry points of the wavefront integrator (wavefront.cu).
#pragma once

#include "renderer.h"

namespace b2pt {

// Trace `spp` samples per pixel starting at frame id r->frame_id + 1
// (OptixRenderer::Render pre-increments frame.id, OptixRenderer.cpp:623).
// accumulate = false: spp must be 1, the frame overwrites r->d_frame
//                     (devicePrograms.cu:704-705);
// accumulate = true : per-pixel sums are added to r->d_accum (K9).
// Asynchronous on r->stream; throws b2pt::Error.
void render_frames(B2ptRenderer* r, int spp, bool accumulate);

// r->d_frame = r->d_accum / r->spp_accumulated (AddPathtracedFrame.frag:18-24).
void resolve_mean(B2ptRenderer* r);

}  // namespace b2pt
