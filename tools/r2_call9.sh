#!/bin/bash
# round 2, GPU call 9: relaxed Philox kernels as the default - the new statistical test alone, then the whole tier
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export B2PT_NO_REBUILD=1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_relaxed_gpu.py -m gpu -q -x > gpurun_out/c9_pytest_relaxed.log 2>&1; echo "pytest relaxed exit $?"; tail -15 gpurun_out/c9_pytest_relaxed.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c9_pytest_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -3 gpurun_out/c9_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/c9_pytest.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/c9_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
