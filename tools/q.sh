mkdir -p gpurun_out/emb
O=gpurun_out/emb
timeout 900 python -m pytest tests/test_tc_parity_gpu.py tests/test_full_size_gpu.py tests/test_edge_cases_gpu.py -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > $O/b_fdiv.json 2> $O/b_fdiv.err; tail -1 $O/b_fdiv.err
