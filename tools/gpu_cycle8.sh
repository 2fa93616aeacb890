#!/bin/bash
# last evidence refresh of round 2 (after the occupancy change of the SIMT tail kernels): tests + the bench lines it moves
mkdir -p gpurun_out/r02h
O=gpurun_out/r02h
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -2 $O/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_C2_1gpu.json 2> $O/bench_C2.err; echo bench rc=$?
timeout 300 python bench.py --workload C1 --steps 20 --warmup 5 --no-cpu > $O/bench_C1.json 2>/dev/null; echo c1 rc=$?
timeout 300 python bench.py --workload C3mpc --steps 5 --warmup 3 > $O/bench_C3mpc_1gpu.json 2>/dev/null; echo c3 rc=$?
timeout 600 python bench.py --workload C4 --steps 1 --warmup 1 > $O/bench_C4_1gpu.json 2>/dev/null; echo c4 rc=$?
timeout 900 python bench.py --workload C5 --steps 2 --warmup 3 --no-cpu > $O/bench_C5_B256.json 2>/dev/null; echo c5 rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/bf16_launches.csv python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu > $O/ncu_bench.log 2>&1; echo ncu list rc=$?
