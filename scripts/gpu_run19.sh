cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_workloads_gpu.py -q -m multigpu --timeout 200 2>&1 | tail -3
timeout 300 $TR --nproc-per-node 2 --master-port 29701 -m cme213_sp18_b200.cli.heat -m --nx 32768 --ny 32768 --iters 20 --no-cpu 2>&1 | tail -3
timeout 300 $TR --nproc-per-node 2 --master-port 29702 -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check 2>&1 | tail -2
timeout 300 python -m cme213_sp18_b200.cli.pagerank 2>&1 | tail -19 | head -6
