set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python scripts/profile_step.py --out gpurun_out/step_wide.json 2>&1 | tail -30
timeout 600 python scripts/profile_step.py --rows 8192 --out gpurun_out/step_wide_8192.json 2>&1 | tail -30
timeout 600 python scripts/profile_step.py --hidden 100 --rows 3200 --iters 20 --out gpurun_out/step_small.json 2>&1 | tail -30
