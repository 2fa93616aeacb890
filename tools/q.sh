mkdir -p gpurun_out/c4
O=gpurun_out/c4
timeout 600 python -m pytest tests/test_replay.py -x -q 2>&1 | tail -5
timeout 900 python bench.py --workload C4 --steps 1 --warmup 1 > $O/bench_C4_1gpu.json 2> $O/c4.err; echo rc=$?; tail -3 $O/c4.err
GNNUDS_C4_PER_EVENT=1 timeout 900 python bench.py --workload C4 --steps 1 --warmup 1 > $O/bench_C4_1gpu_per_event.json 2>/dev/null
