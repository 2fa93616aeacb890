#!/bin/bash
# round 2, cycle 14: launch list of one 12-chunk gradient of ONE decision vector (the second of two calls)
mkdir -p gpurun_out/g1
GNNUDS_GRAPHS=0 python tools/one_grad.py 1 > gpurun_out/g1/plain.log 2>&1; tail -1 gpurun_out/g1/plain.log
GNNUDS_GRAPHS=0 timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 2039 -c 2039 --csv \
  --log-file gpurun_out/g1/grad1_launches.csv python tools/one_grad.py 1 > gpurun_out/g1/ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(l for l in open('gpurun_out/g1/grad1_launches.csv') if l.startswith('"'))]
hd=rows[0]; ik=hd.index('Kernel Name'); iv=hd.index('Metric Value'); iu=hd.index('Metric Unit')
t=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    v=float(r[iv].replace(',','')); v = v/1e3 if r[iu] in ('ns','nsecond') else v
    k=r[ik].split('(')[0]; t[k][0]+=1; t[k][1]+=v
tot=sum(v[1] for v in t.values())
print('total us',round(tot), 'launches', sum(v[0] for v in t.values()))
for k,v in sorted(t.items(), key=lambda kv:-kv[1][1])[:25]: print('%-60s %5d %9.0f us %5.1f%%  avg %.1f us'%(k[:60],v[0],v[1],100*v[1]/tot, v[1]/v[0]))
PY
