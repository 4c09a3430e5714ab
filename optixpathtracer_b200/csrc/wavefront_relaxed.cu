// wavefront_relaxed.cu — the Philox instantiation of the wavefront kernels (K4, K6, K7a-c, K9) in relaxed arithmetic:
// FMA contraction, approximate division / square root, hardware sin / cos / exp (build.py: --use_fast_math).
// This is what rng_mode PHILOX — the production generator — runs; B2PT_RELAXED=0 sends it through the exact kernels
// of wavefront.cu instead (the A/B: c2 +10 %, c3 +9 %, the conductor / dielectric shade kernels twice as fast,
// profiles/r02_ab_relaxed.txt).  The LCG modes — the ones that reproduce the reference bit for bit — never run
// through this file.  Checked statistically: tests/test_relaxed_gpu.py (against the exact kernels on the same random
// numbers) and tests/test_image_gpu.py (against the reference's render, noise floor).
#define B2PT_RELAXED_MATH 1
#include "wavefront.cu"
