#!/bin/bash
# Round profile refresh: bench lines (bf16 default incl. CPU leg, fp32, C1, C5, reference arm), the ncu launch list of
# the bench command, and one ncu --set full capture (raw page as CSV) of the dominant kernels.
mkdir -p gpurun_out/prof
O=gpurun_out/prof
python bench.py --steps 3 --warmup 3 > $O/bench_bf16_B4096.json 2> $O/bench_bf16.err; echo bf16 rc=$?
python bench.py --steps 2 --warmup 3 --precision fp32 --no-cpu > $O/bench_fp32_B4096.json 2> $O/bench_fp32.err; echo fp32 rc=$?
python bench.py --steps 3 --warmup 3 --workload C1 --no-cpu > $O/bench_C1_B64.json 2> $O/bench_C1.err; echo C1 rc=$?
python bench.py --steps 2 --warmup 3 --workload C5 --batch 16 --no-cpu > $O/bench_C5_B16.json 2> $O/bench_C5.err; echo C5 rc=$?
python bench.py --steps 3 --warmup 3 --workload C3a --no-cpu > $O/bench_C3a_B8192.json 2> $O/bench_C3a.err; echo C3a rc=$?
python bench.py --steps 3 --warmup 3 --workload C3b --no-cpu > $O/bench_C3b_B8192.json 2> $O/bench_C3b.err; echo C3b rc=$?
python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_reference_arm.json 2> $O/bench_ref.err; echo ref rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/bf16_launches.csv python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu > $O/ncu_bench.log 2>&1; echo ncu list rc=$?
timeout 900 ncu --set full --clock-control none -k regex:"sp_conv_kernel|tc_rows_kernel|agg_large" -s 60 -c 20 -o /tmp/capf -f python tools/one_forward.py C2 512 bf16 > $O/ncu_full.log 2>&1; echo full rc=$?
ncu -i /tmp/capf.ncu-rep --page raw --csv > $O/full_raw.csv 2>/dev/null
ls -la $O
