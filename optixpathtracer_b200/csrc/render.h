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

// image_out.cu (K10): display transform + file writers on the running mean
void display_to_host(B2ptRenderer* r, bool apply_oetf, bool top_first, uint8_t* h_rgb8);
void write_bmp(B2ptRenderer* r, const char* path, bool apply_oetf);
void write_png(B2ptRenderer* r, const char* path, bool apply_oetf);
void write_exr(B2ptRenderer* r, const char* path);

}  // namespace b2pt
