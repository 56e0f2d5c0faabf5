cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
OUT=$GRAFT_REPO_ROOT/gpurun_out/hw_baselines.log
: > $OUT
echo "=== reference hw3 (unmodified, sm_100): 4096^2 order 8, 100 iters" >> $OUT
(cd baseline/_ref/hw3code && printf "4096 4096\n1 1\n100\n8\n" > params.in && timeout 300 ./main -gb >> $OUT 2>&1)
echo "=== ours hw3: same grid" >> $OUT
timeout 300 python -m cme213_sp18_b200.cli.heat -g -b -s --iters 100 >> $OUT 2>&1
echo "=== ours hw3: 32768^2 order 8, 20 iters, one GPU, no CPU oracle" >> $OUT
timeout 300 python -m cme213_sp18_b200.cli.heat -g -b -s --nx 32768 --ny 32768 --iters 20 --no-cpu >> $OUT 2>&1
echo "=== reference hw3: 32768^2 order 8 would need its CPU loop (skipped); 8192^2 20 iters" >> $OUT
(cd baseline/_ref/hw3code && printf "8192 8192\n1 1\n20\n8\n" > params.in && timeout 300 ./main -gb >> $OUT 2>&1)
timeout 300 python -m cme213_sp18_b200.cli.heat -g -b -s --nx 8192 --ny 8192 --iters 20 >> $OUT 2>&1
echo "=== reference hw2 q1 shift (8 doublings)" >> $OUT
(cd baseline/_ref/hw2code && timeout 300 ./main_q1 8 >> $OUT 2>&1)
echo "=== ours hw2 q1 shift (8 doublings; synthetic text of the same size)" >> $OUT
timeout 300 python -m cme213_sp18_b200.cli.shift 8 >> $OUT 2>&1
echo "=== reference hw2 q2 pagerank sweep" >> $OUT
(cd baseline/_ref/hw2code && timeout 600 ./main_q2 >> $OUT 2>&1)
echo "=== ours hw2 q2 pagerank sweep" >> $OUT
timeout 600 python -m cme213_sp18_b200.cli.pagerank >> $OUT 2>&1
echo "=== ours pagerank 64M nodes avg 8, one GPU" >> $OUT
timeout 600 python -m cme213_sp18_b200.cli.pagerank --nodes 67108864 --avg 8 --no-check >> $OUT 2>&1
tail -5 $OUT
