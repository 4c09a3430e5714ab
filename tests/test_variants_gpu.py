"""GPU tier: every A/B build variant of the traversal kernels gives the same bits.

The switches (B2PT_STEP, B2PT_SSTACK, B2PT_BVH4, B2PT_DEFER, B2PT_OCC, B2PT_TRAVERSAL, B2PT_SMALL_CHUNKS, B2PT_COLLAPSE, B2PT_LEAF
(triangles per leaf of the wide tree), B2PT_SORT_SHADOW (K8 takes its rays from a light-ordered list), and the
shading-side B2PT_CULL_NEE, B2PT_ENGINE_REFILL, B2PT_ENGINE_OCC — how the layered walk engine hands tasks to lanes) are
read once per process, so each variant runs in its own subprocess (tests/variant_probe.py): closest hits must
equal the fp64 brute-force oracle bit for bit, and the occlusion answers, a frame in the reference's random
sequence and an accumulated Philox image must be identical across variants — and across wave capacities.  The
same for the alpha cut-out scene, whose any-hit predicate runs inside the traversal kernels."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

VARIANTS = {
    "default": {},
    "step5_128_byte_nodes": {"B2PT_STEP": "5"},
    "shared_stack_24_levels": {"B2PT_SSTACK": "24"},
    "shared_stack_32_levels": {"B2PT_SSTACK": "32"},
    "binary_nodes_deferred": {"B2PT_BVH4": "0"},
    "binary_nodes_fp64_in_kernel": {"B2PT_BVH4": "0", "B2PT_DEFER": "0"},
    "binary_nodes_occupancy_6": {"B2PT_BVH4": "0", "B2PT_OCC": "6"},
    "thread_per_ray": {"B2PT_TRAVERSAL": "0"},
    "fixed_collapse_big_chunks": {"B2PT_COLLAPSE": "fixed", "B2PT_SMALL_CHUNKS": "0"},
    "walk_engine_refills_every_idle_lane": {"B2PT_ENGINE_REFILL": "1", "B2PT_ENGINE_OCC": "8"},
    "walk_engine_refills_whole_warps": {"B2PT_ENGINE_REFILL": "32", "B2PT_ENGINE_OCC": "4"},
    "every_shadow_ray_traced": {"B2PT_CULL_NEE": "0"},
    "any_hit_nearest_first_break_20": {"B2PT_ANY_UNSORTED": "0", "B2PT_BREAK": "20"},
    "one_triangle_per_leaf": {"B2PT_LEAF": "1"},
    "four_triangles_per_leaf": {"B2PT_LEAF": "4"},
    "eight_triangles_per_leaf_step5": {"B2PT_LEAF": "8", "B2PT_STEP": "5"},
    "shadow_rays_from_a_light_ordered_list": {"B2PT_SORT_SHADOW": "1"},
    "light_ordered_list_serial_layered": {"B2PT_SORT_SHADOW": "1", "B2PT_LAYERED_ENGINE": "0"},
}
# variants whose Philox images are another (equally valid) estimate: everything but `accum_*` must still be equal
NOT_BIT_EQUAL_PHILOX = {"light_ordered_list_serial_layered"}


def _probe(env_extra):
    env = dict(os.environ)
    for k in list(env):
        if k.startswith("B2PT_"):
            del env[k]
    env.update(env_extra)
    p = subprocess.run([sys.executable, str(ROOT / "tests" / "variant_probe.py")], env=env, capture_output=True, text=True,
                       timeout=600, cwd=str(ROOT))
    assert p.returncode == 0, p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("VARIANT ")][-1]
    return json.loads(line[len("VARIANT "):])


@pytest.fixture(scope="module")
def default_probe():
    return _probe({})


@pytest.mark.parametrize("name", list(VARIANTS))
def test_variant_gives_the_same_bits(name, default_probe):
    got = default_probe if name == "default" else _probe(VARIANTS[name])
    assert got["closest_equals_oracle"], "closest hits differ from the fp64 brute force"
    for key in ("any", "frame", "accum_auto", "closest_textured", "any_textured", "frame_textured"):
        if key == "accum_auto" and name in NOT_BIT_EQUAL_PHILOX:
            continue   # the serial layered kernel draws its Philox numbers in another order than the walk engine
        assert got[key] == default_probe[key], f"{key} differs from the default build"
    assert got["accum_small_waves"] == got["accum_auto"], "the image depends on the wave capacity"
    if name == "default":
        assert got["node_bytes"] == 96
    if name == "step5_128_byte_nodes":
        assert got["node_bytes"] == 128
