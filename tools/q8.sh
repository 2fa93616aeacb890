mkdir -p gpurun_out/fin8
O=gpurun_out/fin8
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > $O/bench_C2_8gpu.json 2> $O/b2.err; echo rc=$?; tail -1 $O/b2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --workload C3mpc --steps 5 --warmup 3 > $O/bench_C3mpc_8gpu.json 2> $O/b3.err; echo rc=$?; tail -1 $O/b3.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --workload C4 --steps 1 --warmup 1 > $O/bench_C4_8gpu.json 2> $O/b4.err; echo rc=$?; tail -1 $O/b4.err
