mkdir -p gpurun_out/c5
O=gpurun_out/c5
timeout 900 python bench.py --workload C5 --batch 32 --steps 2 --warmup 3 --no-cpu > $O/mb4.json 2> $O/mb4.err; tail -1 $O/mb4.err
GNNUDS_AGG_MINB=5 timeout 900 python bench.py --workload C5 --batch 32 --steps 2 --warmup 3 --no-cpu > $O/mb5.json 2> $O/mb5.err; tail -1 $O/mb5.err
