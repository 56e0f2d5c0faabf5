cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu --timeout 120 -x -k "heat or pagerank" 2>&1 | tail -4
timeout 300 python -m cme213_sp18_b200.cli.heat -g -b -s --iters 100 --no-cpu 2>&1 | tail -4
timeout 300 python -m cme213_sp18_b200.cli.heat -g -b -s --nx 8192 --ny 8192 --iters 20 --no-cpu 2>&1 | tail -4
timeout 300 python -m cme213_sp18_b200.cli.heat -s --nx 32768 --ny 32768 --iters 20 --no-cpu 2>&1 | tail -2
timeout 300 python -m cme213_sp18_b200.cli.heat -s --nx 8192 --ny 8192 --iters 20 --order 2 --no-cpu 2>&1 | tail -2
timeout 300 python -m cme213_sp18_b200.cli.pagerank --no-check 2>&1 | head -5
