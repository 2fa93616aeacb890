#!/bin/bash
# round 2, cycle 9: kept-tape reverse sweep: gradient tests, then C3grad at 256 / 1 / 1024 decision vectors
mkdir -p gpurun_out/kt
python -m pytest tests/test_grad_gpu.py -x -q -m gpu > gpurun_out/kt/pytest_grad.log 2>&1; echo "grad tests rc=$?"
tail -3 gpurun_out/kt/pytest_grad.log
for B in 256 1 1024; do
  python bench.py --workload C3grad --batch $B --steps 5 --warmup 3 > gpurun_out/kt/grad_$B.json 2> gpurun_out/kt/grad_$B.err; echo "B=$B rc=$?"
  GNNUDS_TAPE_MB=0 python bench.py --workload C3grad --batch $B --steps 5 --warmup 3 > gpurun_out/kt/grad_${B}_ckpt.json 2> gpurun_out/kt/grad_${B}_ckpt.err; echo "B=$B ckpt rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/kt/grad_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],2), 'fwd', round(d.get('forward_only_ms',0),2), d.get('grad_rel_err_vs_oracle'))
    except Exception as e: print(f, 'ERR', e)
PY
