"""CPU tier: the host side of bench.py's contract — no GPU, no timing.

* both arms (b2pt and --impl reference) print the SAME `config` object for the same flags (the driver compares them);
* strong scaling splits the step's samples over the GPUs, weak scaling does not;
* profiles/traffic.json (the ncu figures the bench line quotes) parses and holds what bench.py reads from it;
* the default workload is c4, the largest single-GPU configuration of BASELINE.json."""
import argparse
import json
from pathlib import Path

import pytest

import bench

ROOT = Path(__file__).resolve().parent.parent


def _args(**kw):
    d = dict(gpus=1, steps=8, warmup=3, impl="b2pt", config="c4", scaling="weak", spp_per_step=64, wave_log2=0,
             cpu_seconds=None, no_cpu_baseline=False)
    d.update(kw)
    return argparse.Namespace(**d)


def test_default_workload_is_the_largest_single_gpu_config():
    label, spec, w, h, spp, depth, mode = bench.CONFIGS["c4"]
    assert (w, h, spp, depth) == (1920, 1080, 1024, 8)
    biggest = max((c for c in ("c1", "c2", "c3", "c4")), key=lambda c: bench.CONFIGS[c][2] * bench.CONFIGS[c][3] * bench.CONFIGS[c][4])
    assert biggest == "c4"


@pytest.mark.parametrize("cid", ["c1", "c4"])
def test_both_arms_print_the_same_config(cid):
    cfg = bench.CONFIGS[cid]
    scene = bench.make_scene(cfg[1])
    a = bench.workload_config(_args(config=cid), cfg, scene)
    b = bench.workload_config(_args(config=cid, impl="reference", steps=2, warmup=1), cfg, scene)
    assert a == b and a["config_id"] == cid and "model" not in a
    json.dumps(a)


def test_strong_scaling_splits_the_step():
    assert bench.spp_per_gpu(_args(gpus=8, scaling="strong")) == 8
    assert bench.spp_per_gpu(_args(gpus=8, scaling="weak")) == 64
    with pytest.raises(AssertionError):
        bench.spp_per_gpu(_args(gpus=3, scaling="strong"))


def test_traffic_json_holds_what_the_bench_line_quotes():
    d = json.load(open(ROOT / "profiles" / "traffic.json"))
    for key in ("k_layered_engine_issue_active", "k_layered_engine_lanes_per_inst", "k_layered_engine_thread_inst_per_task",
                "k_layered_engine_dram_bytes_per_task"):
        assert d["c4"][key] > 0
    for cid in ("c2", "c5"):
        assert d[cid]["k_closest_dram_bytes_per_ray"] > 0
    assert 0.0 < d["k_closest_limits_readme_scene"]["incoherent"]["issue_slots_busy"] < 1.0
