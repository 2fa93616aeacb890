#!/bin/bash
# round 2, cycle 13: skinny fp32 GEMMs (forward + VJP): gradient tests, crossover of GNNUDS_SKINNY_ROWS
mkdir -p gpurun_out/sk2
python -m pytest tests/test_grad_gpu.py tests/test_edge_cases_gpu.py tests/test_parity_gpu.py -q -m gpu -x > gpurun_out/sk2/pytest.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/sk2/pytest.log
run() {  # threshold, batch
  GNNUDS_SKINNY_ROWS=$1 python bench.py --workload C3grad --batch $2 --steps 5 --warmup 3 2> gpurun_out/sk2/grad_$2_$1.err | tee gpurun_out/sk2/grad_$2_$1.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('skinny<=$1 grad B=$2', round(d['ms_per_step'],2), 'fwd', round(d['forward_only_ms'],2), d['grad_rel_err_vs_oracle'])"
}
run 0 64
for S in 8192 65536; do for B in 1 16 64 256; do run $S $B; done; done
GNNUDS_SKINNY_ROWS=65536 python tools/roll_latency.py fp32
