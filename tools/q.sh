mkdir -p gpurun_out/par
O=gpurun_out/par
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -4 $O/pytest_gpu.log
timeout 300 python bench.py --workload C1 --steps 20 --warmup 5 --no-cpu > $O/bench_C1.json 2> $O/bench_C1.err; echo c1 rc=$?
GNNUDS_SIDE_PAR=0 timeout 300 python bench.py --workload C1 --steps 20 --warmup 5 --no-cpu > $O/bench_C1_serial.json 2>/dev/null
timeout 300 python bench.py --workload C3mpc --steps 5 --warmup 3 --batch 1024 > $O/bench_C3mpc_1024.json 2> $O/bench_C3mpc.err; echo c3 rc=$?
GNNUDS_SIDE_PAR=0 timeout 300 python bench.py --workload C3mpc --steps 5 --warmup 3 --batch 1024 > $O/bench_C3mpc_1024_serial.json 2>/dev/null
for B in 1 32; do timeout 200 python bench.py --workload C3grad --batch $B --steps 5 --warmup 3 > $O/b_grad$B.json 2> $O/b_grad$B.err; tail -2 $O/b_grad$B.err; done
