#!/bin/bash
# One measure cycle on the GPU box: TC parity diag (3 cases), bench (no CPU leg), ncu launch list, one full capture.
mkdir -p gpurun_out
timeout 300 python tools/tc_diag.py c2_shq_gat_conv_10 c1_ast_gcn_gru ast_gat_gru_noflood > gpurun_out/tc_diag.log 2>&1; echo "rc=$?" >> gpurun_out/tc_diag.log
grep -E "==|predict|rc=" gpurun_out/tc_diag.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo bench rc=$?
cut -c1-330 gpurun_out/bench_bf16.json; tail -3 gpurun_out/bench_bf16.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_bf16.csv python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo ncu rc=$?
if [ "$1" = "full" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k tc_rows_kernel -c 14 -o gpurun_out/tc_rows_cur -f python tools/one_forward.py C2 512 bf16 > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
fi
