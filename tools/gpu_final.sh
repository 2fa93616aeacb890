#!/bin/bash
# end of round: the GPU test suite, smoke(), the default bench line and the reference arm
mkdir -p gpurun_out/final3
python -m pytest tests -q -m gpu > gpurun_out/final3/pytest_gpu.log 2>&1; echo "gpu tests rc=$?"
tail -3 gpurun_out/final3/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final3/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/final3/smoke.log
python bench.py > gpurun_out/final3/bench_default.json 2> gpurun_out/final3/bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final3/bench_reference.json 2> gpurun_out/final3/bench_reference.err; echo "ref rc=$?"
python - <<'PY'
import json
for f in ('bench_default','bench_reference'):
    try:
        d=json.loads(open('gpurun_out/final3/%s.json'%f).read().strip().splitlines()[-1])
        print(f, d.get('value'), d.get('ms_per_step'), (d.get('e2e') or {}).get('value'), (d.get('roofline') or {}).get('frac'), d.get('clocks'))
    except Exception as e: print(f,'ERR',e)
PY
