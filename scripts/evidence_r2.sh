#!/bin/bash
# Round-2 evidence on ONE GPU: -g 4 GEMM table, hw1/hw4 timings, and ncu captures of the fused
# tensor-core kernels (forward with layer-2 riders; dW1 with the update warps; backward tail).
#   gpurun --timeout 900 -- 'bash scripts/evidence_r2.sh'
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
O=gpurun_out/evidence_r2; mkdir -p $O
python -m baseline.build_homeworks > $O/build_homeworks.log 2>&1
timeout 300 python -m cme213_sp18_b200.cli.gemm_bench > $O/gemm_bench.log 2>&1; tail -14 $O/gemm_bench.log
timeout 300 python scripts/time_hw14.py --out $O/hw14_timing.json > $O/hw14.log 2>&1; tail -3 $O/hw14.log
BENCH="python bench.py --steps 3 --warmup 2 --no-e2e --no-check --no-graph"
$BENCH > $O/plain.log 2>&1 || { echo "plain bench failed"; tail -5 $O/plain.log; exit 1; }
NCU="ncu --set full --clock-control none --import-source on -f"
$NCU -k regex:fwd_rider_kernel -s 2 -c 1 -o $O/prof_fwd_rider $BENCH > $O/ncu_fwd.log 2>&1
$NCU -k regex:gemm_mnmajor2_kernel -s 2 -c 1 -o $O/prof_dw1_update $BENCH > $O/ncu_dw1.log 2>&1
$NCU -k regex:bwd_tail_kernel -s 2 -c 1 -o $O/prof_bwd_tail $BENCH > $O/ncu_tail.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 30 --csv --log-file $O/launches_wide_1gpu.csv $BENCH > $O/ncu_l.log 2>&1
ls $O
