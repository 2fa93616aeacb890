#!/bin/bash
# tests + the MPC / replay workloads
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 600 python bench.py --workload C3mpc --steps 5 --warmup 3 > gpurun_out/r02_bench_C3mpc_1gpu.json 2> gpurun_out/c3mpc.err; echo c3mpc rc=$?; cut -c1-600 gpurun_out/r02_bench_C3mpc_1gpu.json; tail -3 gpurun_out/c3mpc.err
timeout 900 python bench.py --workload C4 --steps 1 --warmup 1 > gpurun_out/r02_bench_C4_1gpu.json 2> gpurun_out/c4.err; echo c4 rc=$?; cut -c1-900 gpurun_out/r02_bench_C4_1gpu.json; tail -3 gpurun_out/c4.err
