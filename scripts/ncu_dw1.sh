cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python scripts/profile_step.py --iters 1"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_mnmajor_kernel -s 3 -c 1 -f -o gpurun_out/prof_dw1 $CMD > gpurun_out/ncu_dw1.log 2>&1
tail -3 gpurun_out/ncu_dw1.log
