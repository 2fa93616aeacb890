#!/bin/bash
# round 2, cycle 10: kept-tape reverse sweep with balanced blocks; launch list of one 12-chunk gradient (256 decision vectors)
mkdir -p gpurun_out/kt2
python -m pytest tests/test_grad_gpu.py -x -q -m gpu > gpurun_out/kt2/pytest_grad.log 2>&1; echo "grad tests rc=$?"
tail -2 gpurun_out/kt2/pytest_grad.log
for B in 256 1024 1; do
  python bench.py --workload C3grad --batch $B --steps 5 --warmup 3 > gpurun_out/kt2/grad_$B.json 2> gpurun_out/kt2/grad_$B.err; echo "B=$B rc=$?"
done
GNNUDS_GRAPHS=0 timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 6117 -c 2039 --csv \
  --log-file gpurun_out/kt2/grad_launches.csv python bench.py --workload C3grad --batch 256 --steps 1 --warmup 3 > gpurun_out/kt2/ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import json,glob,csv,collections
for f in sorted(glob.glob('gpurun_out/kt2/grad_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],2), 'fwd', round(d.get('forward_only_ms',0),2), d['gpu_launches']//d['steps'])
    except Exception as e: print(f, 'ERR', e)
try:
    rows=[r for r in csv.reader(l for l in open('gpurun_out/kt2/grad_launches.csv') if l.startswith('"'))]
    hd=rows[0]; ik=hd.index('Kernel Name'); iv=hd.index('Metric Value'); iu=hd.index('Metric Unit')
    t=collections.defaultdict(lambda:[0,0.0])
    for r in rows[1:]:
        v=float(r[iv].replace(',','')); v = v/1e3 if r[iu] in ('ns','nsecond') else v
        k=r[ik].split('(')[0]; t[k][0]+=1; t[k][1]+=v
    tot=sum(v[1] for v in t.values())
    print('total us',round(tot), 'launches', sum(v[0] for v in t.values()))
    for k,v in sorted(t.items(), key=lambda kv:-kv[1][1])[:25]: print('%-60s %5d %9.0f us %5.1f%%'%(k[:60],v[0],v[1],100*v[1]/tot))
except Exception as e: print('csv ERR', e)
PY
