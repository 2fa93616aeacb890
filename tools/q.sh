mkdir -p gpurun_out/fin
O=gpurun_out/fin
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_C2_2gpu.json 2> $O/b2.err; echo rc=$?; tail -2 $O/b2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload C3mpc --steps 5 --warmup 3 > $O/bench_C3mpc_2gpu.json 2> $O/b3.err; echo rc=$?; tail -2 $O/b3.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --impl reference --steps 1 --warmup 0 > $O/bench_ref_2gpu.json 2> $O/b4.err; echo rc=$?
