#!/bin/bash
# One plain run, then ncu captures (full set, one instance each) of the hot MLP kernels and a
# launch list of a short bench run.  Reports land in gpurun_out/.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python scripts/profile_step.py --iters 1"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
NCU="ncu --set full --clock-control none --import-source on -f"
$NCU -k regex:'gemm_kmajor_kernel<1>' -s 1 -c 1 -o gpurun_out/prof_fwd1 $CMD > gpurun_out/ncu_fwd1.log 2>&1
$NCU -k regex:'gemm_kmajor_kernel<2>' -s 1 -c 1 -o gpurun_out/prof_fwd2 $CMD > gpurun_out/ncu_fwd2.log 2>&1
$NCU -k regex:gemm_mnmajor_kernel -s 4 -c 1 -o gpurun_out/prof_dw1 $CMD > gpurun_out/ncu_dw1.log 2>&1
$NCU -k regex:gemm_mnmajor_kernel -s 1 -c 1 -o gpurun_out/prof_dw2 $CMD > gpurun_out/ncu_dw2.log 2>&1
$NCU -k regex:bwd_dz1_kernel -s 1 -c 1 -o gpurun_out/prof_dz1 $CMD > gpurun_out/ncu_dz1.log 2>&1
$NCU -k regex:allreduce_sgd -s 1 -c 1 -o gpurun_out/prof_update $CMD > gpurun_out/ncu_update.log 2>&1
BENCH="python bench.py --steps 3 --warmup 3 --no-e2e --no-graph"
$BENCH > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
ls -la gpurun_out | grep -E "prof_|launches"
