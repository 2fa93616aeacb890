#!/bin/bash
# multi-GPU lines: tools/gpu_multi.sh N   (C2 weak scaling, C3mpc and C4 strong scaling at N ranks)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --workload C3mpc --steps 5 --warmup 3 > gpurun_out/r02_bench_C3mpc_${N}gpu.json 2> gpurun_out/c3mpc_$N.err; echo c3mpc rc=$?; cut -c1-300 gpurun_out/r02_bench_C3mpc_${N}gpu.json; grep -o '"allgather_us": [0-9.]*' gpurun_out/r02_bench_C3mpc_${N}gpu.json
timeout 900 $TR bench.py --gpus $N --workload C4 --steps 1 --warmup 1 > gpurun_out/r02_bench_C4_${N}gpu.json 2> gpurun_out/c4_$N.err; echo c4 rc=$?; cut -c1-300 gpurun_out/r02_bench_C4_${N}gpu.json
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_C2_${N}gpu.json 2> gpurun_out/c2_$N.err; echo c2 rc=$?; cut -c1-300 gpurun_out/r02_bench_C2_${N}gpu.json
tail -3 gpurun_out/c3mpc_$N.err gpurun_out/c4_$N.err gpurun_out/c2_$N.err | grep -v "^$" | tail -8
