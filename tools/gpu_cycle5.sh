#!/bin/bash
# round-2 profile set: C5 at its configured B=256, launch list, ncu full captures of the hot kernels reduced to CSV on the box
mkdir -p gpurun_out
if [ "$1" != "ncu" ]; then
timeout 900 python bench.py --workload C5 --steps 2 --warmup 3 --no-cpu > gpurun_out/r02_bench_C5_B256.json 2> gpurun_out/c5.err; echo c5 rc=$?; tail -2 gpurun_out/c5.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_bf16_launches.csv python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo ncu rc=$?
fi
timeout 900 ncu --set full --clock-control none -k regex:"sp_agg_kernel|tc_rows_kernel" -c 12 -o /tmp/r02_full -f python tools/one_forward.py C2 256 bf16 > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
ncu -i /tmp/r02_full.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_raw.csv 2> gpurun_out/ncu_export.err; echo export rc=$?
ls -la /tmp/r02_full.ncu-rep gpurun_out/ | head -20
