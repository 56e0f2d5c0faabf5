#!/bin/bash
# Native runtime vs Python feeder on one B200: tests, then end-to-end numbers for both configs.
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_mlp_gpu.py -q -m "gpu and not multigpu" --timeout 300 -x 2>&1 | tail -5
for f in native python; do
  timeout 300 python bench.py --config small --steps 400 --warmup 20 --feeder $f 2>&1 | tail -1 | tee gpurun_out/feeder_small_$f.json
  timeout 300 python bench.py --steps 20 --warmup 5 --feeder $f 2>&1 | tail -1 | tee gpurun_out/feeder_wide_$f.json
done
