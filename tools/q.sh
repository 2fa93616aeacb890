mkdir -p gpurun_out/emb
O=gpurun_out/emb
timeout 900 python -m pytest tests/test_tc_parity_gpu.py tests/test_parity_gpu.py tests/test_full_size_gpu.py tests/test_edge_cases_gpu.py tests/test_large_graph_gpu.py -x -q 2>&1 | tail -4
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu > $O/b_mma.json 2> $O/b_mma.err; tail -1 $O/b_mma.err
GNNUDS_EMBED_MMA=0 timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu > $O/b_simt.json 2>/dev/null
