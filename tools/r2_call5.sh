#!/bin/bash
# round 2, GPU call 5: no-alloc entry point test (alone, then after everything), full GPU tier with the alpha predicate in the
# wide kernels, bench lines of every config + reference arm, textured vs untextured, shared-stack A/B, launch list, ncu captures
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c5_pytest_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -5 gpurun_out/c5_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/c5_pytest.log
timeout 900 python bench.py > gpurun_out/c5_bench_c4.json 2> gpurun_out/c5_bench_c4.err; echo "bench c4 exit $?"; tail -3 gpurun_out/c5_bench_c4.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c5_bench_ref_c4.json 2> gpurun_out/c5_bench_ref.err; echo "ref exit $?"; cut -c1-300 gpurun_out/c5_bench_ref_c4.json
for c in c1 c2 c3 c5 tex auto; do
  timeout 600 python bench.py --config $c --steps 4 --warmup 3 > gpurun_out/c5_bench_$c.json 2> gpurun_out/c5_bench_$c.err; echo "bench $c exit $?"
done
python - <<'PY'
import json
for c in ("c4", "c1", "c2", "c3", "c5", "tex", "auto"):
    try:
        d = json.load(open(f"gpurun_out/c5_bench_{c}.json")); r = d["roofline"]; k = r.get("traversal", r)
        print(c, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), {a: round(b, 3) for a, b in r["share_of_step"].items()},
              "K5 ms/launch", round(k["ms_per_launch"], 3), "nodes/ray", round(k["nodes_per_ray"], 2), "build ms", d["device"]["bvh_build_ms"],
              "wide nodes", d["device"]["bvh_wide_nodes"], "cpu", d.get("cpu_baseline", {}).get("value"))
    except Exception as e:
        print(c, "unreadable:", e)
PY
export AB_ARGS="--config c2 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_SSTACK=16" "B2PT_SSTACK=24" "B2PT_SSTACK=32" 2>&1 | tee gpurun_out/c5_ab_sstack_c2.txt
export AB_ARGS="--config c5 --steps 3 --warmup 2"
bash tools/ab.sh "B2PT_SSTACK=16" "B2PT_SSTACK=24" "B2PT_SSTACK=32" 2>&1 | tee gpurun_out/c5_ab_sstack_c5.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/c5_launches_c4.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c5_ncu_launches.log 2>&1; echo "ncu launches exit $?"
for k in k_layered_engine k_closest_w k_any_w k_surface k_layered_prep k_layered_finish k_raygen k_accumulate; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/c5_$k python bench.py --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c5_ncu_$k.log 2>&1; echo "ncu $k exit $?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_closest_w -c 1 -o gpurun_out/c5_tex_k_closest_w python bench.py --config tex --steps 1 --warmup 0 --spp-per-step 16 --no-cpu-baseline > gpurun_out/c5_ncu_tex.log 2>&1; echo "ncu tex exit $?"
