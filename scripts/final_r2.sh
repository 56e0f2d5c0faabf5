#!/bin/bash
# Round-2 closing measurements on N GPUs of one box:   gpurun --gpus N -- 'bash scripts/final_r2.sh N'
# Everything lands in gpurun_out/final/ (copied to profiles/final_r2/ afterwards).
N=${1:-1}
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
O=gpurun_out/final
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() {  # name port args...: one bench.py run on N GPUs, JSON line to $O/name.json
  local name=$1 port=$2; shift 2
  if [ "$N" = 1 ]; then timeout 400 python bench.py --gpus 1 "$@" > $O/$name.log 2>&1
  else timeout 400 $TR --master-port $port bench.py --gpus $N "$@" > $O/$name.log 2>&1; fi
  tail -1 $O/$name.log > $O/$name.json
  python - "$O/$name.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    c = d.get("self_check") or {}
    print(sys.argv[1].split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "e2e", round((d.get("e2e") or {}).get("ms_per_step", 0), 4),
          "check", c.get("ok"), "exposed", (d.get("communication") or {}).get("exposed_allreduce_ms_per_step"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
if [ "$N" -gt 1 ]; then
  timeout 1200 python -m pytest tests/test_mlp_gpu.py tests/test_workloads_gpu.py -q -m gpu -k "multi_gpu and $N" > $O/tests_$N.log 2>&1
  tail -3 $O/tests_$N.log
fi
run wide_$N 29811 --steps 20 --warmup 5
run small_$N 29812 --config small --steps 200 --warmup 20
run fp64_grade_$N 29813 --precision fp64 --config grade --steps 20 --warmup 5
if [ "$N" -gt 1 ]; then
  run wide_${N}_owner 29814 --steps 20 --warmup 5 --no-e2e --no-check --data-mode owner
  run wide_${N}_owner_noprefetch 29815 --steps 20 --warmup 5 --no-e2e --no-check --data-mode owner --no-prefetch
  timeout 300 $TR --master-port 29816 -m cme213_sp18_b200.cli.heat -m --nx 32768 --ny 32768 --iters 20 --no-cpu 2>&1 | grep -E "Slabs" | tee $O/heat_$N.log
fi
if [ "$N" = 8 ]; then
  run fp64_small_$N 29817 --precision fp64 --config small --steps 20 --warmup 5
  run fp64_wide_$N 29818 --precision fp64 --steps 5 --warmup 3 --no-e2e
  timeout 300 $TR --master-port 29819 -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check 2>&1 | grep -E "^N=" | tee $O/pagerank_$N.log
fi
if [ "$N" = 1 ]; then
  run grade_1 29820 --config grade --steps 50 --warmup 10
  run fp64_small_1 29821 --precision fp64 --config small --steps 20 --warmup 5
  run fp64_wide_1 29822 --precision fp64 --steps 3 --warmup 3 --no-e2e --no-check
  timeout 300 python -m cme213_sp18_b200.cli.gemm_bench > $O/gemm_bench.log 2>&1; tail -12 $O/gemm_bench.log
  timeout 300 python -m cme213_sp18_b200.cli.pagerank > $O/pagerank_sweep_ours.log 2>&1; tail -3 $O/pagerank_sweep_ours.log
fi
