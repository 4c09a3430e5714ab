"""GPU tier, collected last: b2pt_trace_any_noalloc (csrc/noalloc.cu) against b2pt_trace_any.

The entry point has no counterpart in the reference (it was added for callers
that cannot cudaMalloc); its oracle is the existing occlusion hook, which the other GPU tests pin against the fp64
brute force.  Same kernel, same rays: the answers are 0/1 bytes and must be EQUAL, on two calls in a row — the
second with fewer rays and with untraced slots (tmax < 0), so anything the first call left in the shared static
scratch would show.  Calls beyond B2PT_NOALLOC_MAX_RAYS must be refused with a message, not truncated.
"""
import numpy as np
import pytest

from tests import cases
from tests.util import random_rays

pytestmark = pytest.mark.gpu
f32 = np.float32


def _rays(scene, rng, n):
    from optixpathtracer_b200 import scenes
    soup, _ = scenes.scene_triangle_soup(scene)
    lo, hi = soup.reshape(-1, 3).min(0), soup.reshape(-1, 3).max(0)
    o, d = random_rays(rng, n, lo - 0.5, hi + 0.5)
    return o.astype(f32), d.astype(f32)


@pytest.mark.parametrize("name", ["mixed", "textured"])
def test_no_alloc_occlusion_equals_the_allocating_hook_twice_in_a_row(name):
    from optixpathtracer_b200 import OptixRenderer
    sc = cases.frame_scenes()[name]
    rng = np.random.default_rng(2024)
    with OptixRenderer(None, sc, bsdf_mode="auto") as r:
        n_max = r.NOALLOC_MAX_RAYS
        # call 1: the largest call the scratch holds, every slot traced
        o1, d1 = _rays(sc, rng, n_max)
        t1 = rng.uniform(0.5, 6.0, n_max).astype(f32)
        want1 = r.TraceAny(o1, d1, t1)
        got1 = r.TraceAnyNoAlloc(o1, d1, t1)
        assert got1.dtype == np.uint8 and np.array_equal(got1, want1)
        assert 0.02 < want1.mean() < 0.98          # both answers occur
        # call 2, straight after: fewer rays, a third of the slots untraced (tmax < 0 -> 0, whatever call 1 left there)
        n2 = 10007
        o2, d2 = _rays(sc, rng, n2)
        t2 = rng.uniform(0.5, 6.0, n2).astype(f32)
        t2[::3] = -1.0
        want2 = r.TraceAny(o2, d2, t2)
        got2 = r.TraceAnyNoAlloc(o2, d2, t2)
        assert np.array_equal(got2, want2)
        assert not got2[::3].any()
        # and call 1 again: the same bytes as before
        assert np.array_equal(r.TraceAnyNoAlloc(o1, d1, t1), want1)


def test_no_alloc_refuses_calls_beyond_its_static_scratch():
    from optixpathtracer_b200 import OptixRenderer
    from optixpathtracer_b200._lib import B2ptError
    sc = cases.frame_scenes()["mixed"]
    rng = np.random.default_rng(7)
    with OptixRenderer(None, sc, bsdf_mode="auto") as r:
        n = r.NOALLOC_MAX_RAYS + 1
        o, d = _rays(sc, rng, n)
        with pytest.raises(B2ptError) as e:
            r.TraceAnyNoAlloc(o, d, 3.0)
        assert e.value.code == -5 and "B2PT_NOALLOC_MAX_RAYS" in str(e.value)
        assert len(r.TraceAnyNoAlloc(o[:0], d[:0], 3.0)) == 0
        # the renderer is still usable after the refusal
        assert np.array_equal(r.TraceAnyNoAlloc(o[:100], d[:100], 3.0), r.TraceAny(o[:100], d[:100], 3.0))
