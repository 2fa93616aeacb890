#!/bin/bash
# round 2, cycle 15: per-sample aggregation kernels (forward + VJP) with 1024 threads when there are few samples
mkdir -p gpurun_out/ag
python -m pytest tests -q -m gpu -x > gpurun_out/ag/pytest.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/ag/pytest.log
for B in 1 16; do
  python bench.py --workload C3grad --batch $B --steps 5 --warmup 3 2> gpurun_out/ag/grad_$B.err | tee gpurun_out/ag/grad_$B.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('grad B=$B', round(d['ms_per_step'],2), 'fwd', round(d['forward_only_ms'],2), d['grad_rel_err_vs_oracle'])"
done
python tools/roll_latency.py fp32
python tools/roll_latency.py bf16
