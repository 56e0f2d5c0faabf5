#!/bin/bash
# One plain run, then ncu captures (full set, one instance each) of the hot MLP kernels and a
# launch list of a short bench run.  Reports land in gpurun_out/.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python scripts/profile_step.py --iters 1"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
NCU="ncu --set full --clock-control none --import-source on -f"
rm -f gpurun_out/prof_*.ncu-rep
$NCU -k regex:gemm_kmajor2_kernel -s 1 -c 1 -o gpurun_out/prof_fwd1_pair $CMD > gpurun_out/ncu_fwd1.log 2>&1
$NCU -k regex:gemm_kmajor_kernel -s 1 -c 1 -o gpurun_out/prof_fwd2_softmax $CMD > gpurun_out/ncu_fwd2.log 2>&1
$NCU -k regex:gemm_mnmajor2_kernel -s 1 -c 1 -o gpurun_out/prof_dw1_pair $CMD > gpurun_out/ncu_dw1.log 2>&1
$NCU -k regex:bwd_tail_kernel -s 1 -c 1 -o gpurun_out/prof_bwd_tail $CMD > gpurun_out/ncu_tail.log 2>&1
BENCH="python bench.py --steps 3 --warmup 3 --no-e2e --no-graph"
timeout 300 python -m pytest tests/test_mlp_gpu.py -q -m "gpu and not multigpu" --timeout 120 -x -k "streaming or native" 2>&1 | tail -3
$BENCH > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
ls -la gpurun_out | grep -E "prof_|launches"
