#!/bin/bash
# tests, bench, and (with "full") ncu full captures of tc_rows / sp_conv on a small C2 forward
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_cur.json 2> gpurun_out/bench_cur.err; echo bench rc=$?
python - <<PY
import json
d=json.load(open("gpurun_out/bench_cur.json"))
print("value %.4g ms %.2f e2e %.4g launches %d fam %s frac %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], {k:(round(v,1) if isinstance(v,float) else v) for k,v in d["roofline"]["family_ms_per_step"].items() if k!="note"}, d["roofline"]["frac"]))
PY
tail -2 gpurun_out/bench_cur.err
if [ "$1" = "full" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tc_rows_kernel|sp_conv_kernel" -c 40 -o gpurun_out/cur_full -f python tools/one_forward.py C2 256 bf16 > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
fi
