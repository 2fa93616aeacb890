#!/bin/bash
# round 2, cycle 16: programmatic dependent launch on the small-tile GEMMs
mkdir -p gpurun_out/pdl
python -m pytest tests -q -m gpu -x > gpurun_out/pdl/pytest.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/pdl/pytest.log
for P in 1 0; do
  echo "== GNNUDS_PDL=$P"
  for B in 1 16; do
    GNNUDS_PDL=$P python bench.py --workload C3grad --batch $B --steps 8 --warmup 3 2> gpurun_out/pdl/grad_${B}_$P.err | tee gpurun_out/pdl/grad_${B}_$P.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('grad B=$B', round(d['ms_per_step'],2), 'fwd', round(d['forward_only_ms'],2), d['grad_rel_err_vs_oracle'])"
  done
  GNNUDS_PDL=$P python tools/roll_latency.py fp32
done
