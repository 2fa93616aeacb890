mkdir -p gpurun_out/par
O=gpurun_out/par
timeout 600 python -m pytest tests/test_mpc_gpu.py tests/test_grad_gpu.py tests/test_edge_cases_gpu.py -x -q 2>&1 | tail -4
for B in 64 1024 8192; do
timeout 300 python bench.py --workload C3mpc --steps 5 --warmup 3 --batch $B > $O/mpc_$B.json 2> $O/mpc_$B.err; tail -2 $O/mpc_$B.err
GNNUDS_GRAPHS=0 timeout 300 python bench.py --workload C3mpc --steps 5 --warmup 3 --batch $B > $O/mpc_${B}_nograph.json 2>/dev/null
done
