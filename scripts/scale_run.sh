#!/bin/bash
# Round-end style scaling run on an 8-GPU box: ours and the reference arm at N = 1, 2, 4, 8,
# the small config, the secondary metrics and a per-phase breakdown.  Logs -> gpurun_out/scale/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/scale
O=gpurun_out/scale
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
PORT=29900
run() {  # N, extra args..., outfile
  local n=$1; shift; local out=$1; shift
  PORT=$((PORT+1))
  if [ "$n" = "1" ]; then timeout 600 python bench.py --gpus 1 "$@" 2>&1 | tail -1 > $O/$out
  else timeout 600 $TR --nproc-per-node $n --master-port $PORT bench.py --gpus $n "$@" 2>&1 | tail -1 > $O/$out; fi
  cut -c1-220 $O/$out
}
for n in 1 2 4 8; do run $n ours_wide_$n.json --steps 20 --warmup 5; done
run 8 ours_wide_8_breakdown.json --steps 10 --warmup 3 --no-e2e --breakdown
run 8 ours_wide_8_nccl.json --steps 20 --warmup 5 --no-e2e --comm nccl
run 8 ours_wide_8_p2p.json --steps 20 --warmup 5 --no-e2e --comm p2p
run 1 ours_small_1.json --config small --steps 200 --warmup 10
run 8 ours_small_8.json --config small --steps 200 --warmup 10
for n in 1 2 4 8; do run $n ref_wide_$n.json --impl reference --steps 3 --warmup 1; done
run 1 ref_small_1.json --impl reference --config small --steps 20 --warmup 3
run 8 ref_small_8.json --impl reference --config small --steps 20 --warmup 3
for n in 4 8; do bash scripts/bench_secondary.sh $n > $O/secondary_$n.log 2>&1; cat $O/secondary_$n.log; done
