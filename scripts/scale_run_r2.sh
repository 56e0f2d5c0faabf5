#!/bin/bash
# Round-2 scaling run on one 8-GPU box: wide config at 8 / 4 / 2 / 1 GPUs, owner-mode at 8,
# small config at 8 and 1.  Results: gpurun_out/scale_r2/*.json (copy to profiles/scale_r2/).
set -u
OUT=gpurun_out/scale_r2; mkdir -p $OUT
run() { # name nproc args...
  name=$1; n=$2; shift 2
  if [ "$n" = 1 ]; then python bench.py --gpus 1 "$@" > $OUT/$name.json 2> $OUT/$name.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) bench.py --gpus $n "$@" > $OUT/$name.json 2> $OUT/$name.err; fi
  echo "== $name rc=$?"; tail -c 300 $OUT/$name.json | head -c 300; echo
}
run wide_8 8 --steps 20 --warmup 5
run wide_8_owner 8 --steps 20 --warmup 5 --data-mode owner --no-check --no-e2e
run wide_4 4 --steps 20 --warmup 5
run wide_2 2 --steps 20 --warmup 5 --no-check
run wide_1 1 --steps 20 --warmup 5 --no-check
run small_8 8 --config small --steps 200 --warmup 20
run small_1 1 --config small --steps 200 --warmup 20 --no-check
