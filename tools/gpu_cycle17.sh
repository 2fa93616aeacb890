#!/bin/bash
# round 2, cycle 17: programmatic dependent launch for the tensor-core kernels at small grids
mkdir -p gpurun_out/pdl2
python -m pytest tests -q -m gpu -x > gpurun_out/pdl2/pytest.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/pdl2/pytest.log
for P in 1 0; do
  echo "== GNNUDS_PDL=$P"
  GNNUDS_PDL=$P python tools/roll_latency.py bf16
  GNNUDS_PDL=$P python bench.py --workload C1 --batch 64 --steps 20 --warmup 5 --no-cpu 2> gpurun_out/pdl2/c1_$P.err | tee gpurun_out/pdl2/c1_$P.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C1 B=64 ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']/1e6,2))"
  GNNUDS_PDL=$P python bench.py --workload C3mpc --batch 64 --steps 10 --warmup 5 --no-cpu 2> gpurun_out/pdl2/mpc64_$P.err | tee gpurun_out/pdl2/mpc64_$P.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C3mpc pop 64 ms', round(d['ms_per_step'],3))"
done
