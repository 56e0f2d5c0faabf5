#!/bin/bash
# Round-end style check on a B200 box: GPU tests, smoke(), the headline bench.
# usage (from the dev container): gpurun --timeout 900 -- 'bash scripts/gpu_verify.sh'
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
timeout 800 python -m pytest tests -q -m gpu --timeout 300 > gpurun_out/verify_pytest_full.log 2>&1; tail -5 gpurun_out/verify_pytest_full.log | tee gpurun_out/verify_pytest.log; grep -E "^(FAILED|ERROR)" gpurun_out/verify_pytest_full.log | head -10
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" 2>&1 | tail -2 | tee gpurun_out/verify_smoke.log
timeout 300 python bench.py --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/verify_bench_wide.json
