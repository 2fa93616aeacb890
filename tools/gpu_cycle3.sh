#!/bin/bash
# sp_agg_kernel bring-up cycle: TC parity diag under a timeout (a hang must not take the box), TC tests, bench, traces
mkdir -p gpurun_out
timeout 240 python tools/tc_diag.py c2_shq_gat_conv_10 c1_ast_gcn_gru ast_gat_gru_noflood c3_ast_act_5 > gpurun_out/tc_diag.log 2>&1; echo "rc=$?" >> gpurun_out/tc_diag.log
grep -E "==|predict|rc=|rror" gpurun_out/tc_diag.log | head -40
if grep -q "rc=0" gpurun_out/tc_diag.log; then
  timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_cur.json 2> gpurun_out/bench_cur.err; echo bench rc=$?
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_cur.json"))
print("value %.4g ms %.2f e2e %.4g launches %d fam %s frac %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], {k:(round(v,1) if isinstance(v,float) else v) for k,v in d["roofline"]["family_ms_per_step"].items() if k!="note"}, d["roofline"]["frac"]))
PY
  tail -2 gpurun_out/bench_cur.err
  GNNUDS_TRACE_FILE=gpurun_out/sp_agg_trace_x.txt GNNUDS_TRACE_LAUNCH=0 timeout 300 python tools/one_forward.py C2 512 bf16 > /dev/null 2>&1
  GNNUDS_TRACE_FILE=gpurun_out/sp_agg_trace_e.txt GNNUDS_TRACE_LAUNCH=1 timeout 300 python tools/one_forward.py C2 512 bf16 > /dev/null 2>&1
  head -14 gpurun_out/sp_agg_trace_x.txt; sed -n 1p gpurun_out/sp_agg_trace_e.txt; sed -n 20,30p gpurun_out/sp_agg_trace_e.txt
fi
