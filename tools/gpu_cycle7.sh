#!/bin/bash
# end-of-round-2 evidence, final kernels: full GPU test suite, bench lines of every workload, ncu launch list of the bench
# command and one ncu --set full capture of the hot kernels (raw page reduced to CSV on the box)
mkdir -p gpurun_out/r02g
O=gpurun_out/r02g
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -3 $O/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_C2_1gpu.json 2> $O/bench_C2.err; echo bench rc=$?; tail -2 $O/bench_C2.err
timeout 300 python bench.py --workload C1 --steps 20 --warmup 5 --no-cpu > $O/bench_C1.json 2> $O/bench_C1.err; echo c1 rc=$?
timeout 300 python bench.py --workload C3mpc --steps 5 --warmup 3 > $O/bench_C3mpc_1gpu.json 2> $O/bench_C3mpc.err; echo c3 rc=$?
timeout 300 python bench.py --workload C3grad --steps 5 --warmup 3 > $O/bench_C3grad.json 2> $O/bench_C3grad.err; echo c3g rc=$?
timeout 300 python bench.py --workload C3grad --batch 1 --steps 5 --warmup 3 > $O/bench_C3grad_B1.json 2>/dev/null; echo c3g1 rc=$?
timeout 600 python bench.py --workload C4 --steps 1 --warmup 1 > $O/bench_C4_1gpu.json 2> $O/bench_C4.err; echo c4 rc=$?
timeout 900 python bench.py --workload C5 --steps 2 --warmup 3 --no-cpu > $O/bench_C5_B256.json 2> $O/bench_C5.err; echo c5 rc=$?
timeout 300 python bench.py --precision fp32 --steps 2 --warmup 3 --no-cpu > $O/bench_fp32_B4096.json 2> $O/bench_fp32.err; echo fp32 rc=$?
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_reference_arm.json 2> $O/bench_ref.err; echo ref rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/bf16_launches.csv python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu > $O/ncu_bench.log 2>&1; echo ncu list rc=$?
timeout 900 ncu --set full --clock-control none -k regex:"sp_agg|tp_conv_kernel|tc_rows_kernel" -c 16 -o /tmp/r02g_full -f python tools/one_forward.py C2 256 bf16 > $O/ncu_full.log 2>&1; echo full rc=$?
ncu -i /tmp/r02g_full.ncu-rep --page raw --csv > $O/ncu_full_raw.csv 2> $O/ncu_export.err; echo export rc=$?
ls -la $O
