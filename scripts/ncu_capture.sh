#!/bin/bash
# One plain run, then ncu captures of the hot MLP kernels (full set, one instance each) and a launch list.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python scripts/profile_step.py --iters 1"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_kmajor_kernel -s 2 -c 1 -f -o gpurun_out/prof_fwd1 $CMD > gpurun_out/ncu_fwd1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_mnmajor_kernel -s 2 -c 1 -f -o gpurun_out/prof_dw1 $CMD > gpurun_out/ncu_dw1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bwd_dz1_kernel -s 2 -c 1 -f -o gpurun_out/prof_dz1 $CMD > gpurun_out/ncu_dz1.log 2>&1
BENCH="python bench.py --steps 2 --warmup 1 --no-e2e"
$BENCH > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
ls -la gpurun_out | tail -20
