#!/bin/bash
# round 2, cycle 12: one-round 16 x 32 tile fp32 GEMM for small row counts (rows_gemm_f32_skinny)
mkdir -p gpurun_out/sk
python -m pytest tests -q -m gpu -x > gpurun_out/sk/pytest_gpu.log 2>&1; echo "gpu tests rc=$?"
tail -4 gpurun_out/sk/pytest_gpu.log
for S in 0 2048 8192; do
  echo "== GNNUDS_SKINNY_ROWS=$S"
  GNNUDS_SKINNY_ROWS=$S python tools/roll_latency.py fp32
  for B in 1 16; do
    GNNUDS_SKINNY_ROWS=$S python bench.py --workload C3grad --batch $B --steps 5 --warmup 3 2> gpurun_out/sk/grad_${B}_$S.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('grad B=$B', round(d['ms_per_step'],2), 'fwd', round(d['forward_only_ms'],2), d['grad_rel_err_vs_oracle'])"
  done
done
