#!/bin/bash
# sp_agg_kernel cycle: TC parity diag under a timeout (a hang must not take the box), GPU tests, bench, clock64 traces
mkdir -p gpurun_out
timeout 240 python tools/tc_diag.py c2_shq_gat_conv_10 c1_ast_gcn_gru ast_gat_gru_noflood c3_ast_act_5 > gpurun_out/tc_diag.log 2>&1; echo "rc=$?" >> gpurun_out/tc_diag.log
grep -E "==|predict|rc=|rror" gpurun_out/tc_diag.log | head -40
if grep -q "rc=0" gpurun_out/tc_diag.log; then
  if [ "$1" != "notest" ]; then timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6; fi
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_cur.json 2> gpurun_out/bench_cur.err; echo bench rc=$?
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_cur.json"))
print("value %.4g ms %.2f e2e %.4g numpy %.4g launches %d fam %s frac %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"].get("numpy_api",{}).get("value",0), d["gpu_launches"], {k:(round(v,1) if isinstance(v,float) else v) for k,v in d["roofline"]["family_ms_per_step"].items() if k!="note"}, d["roofline"]["frac"]))
PY
  tail -2 gpurun_out/bench_cur.err
  GNNUDS_TRACE_FILE=gpurun_out/sp_agg_trace_x.txt GNNUDS_TRACE_LAUNCH=0 timeout 300 python tools/one_forward.py C2 512 bf16 > /dev/null 2>&1
  GNNUDS_TRACE_FILE=gpurun_out/sp_agg_trace_e.txt GNNUDS_TRACE_LAUNCH=1 timeout 300 python tools/one_forward.py C2 512 bf16 > /dev/null 2>&1
  python tools/sa_trace.py gpurun_out/sp_agg_trace_x.txt; python tools/sa_trace.py gpurun_out/sp_agg_trace_e.txt
fi
