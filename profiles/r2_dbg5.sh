#!/bin/bash
# launch list of one HV encode (lena 512x512): where the time goes, level by level
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_hv_launches.csv python profiles/hv_once.py 1 > gpurun_out/ncu_hv.log 2>&1; tail -1 gpurun_out/ncu_hv.log
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/r2_hv_launches.csv')) if len(r)>10]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); ui=h.index('Metric Unit'); gi=h.index('Grid Size')
agg=collections.defaultdict(lambda:[0,0.0]); big=[]
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v = v/1e3 if u=='ns' else (v if u in ('us','usecond') else v*1e3)
    k=r[ki].split('(')[0][-40:]; agg[k][0]+=1; agg[k][1]+=v
    if 'search' in k: big.append((v, r[gi]))
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:8]: print('%-42s %4d %10.1f us'%(k,n,t))
big.sort(reverse=True)
print('largest search launches (us, grid):', [(round(a),b) for a,b in big[:14]])
print('search launches under 200 us:', sum(1 for a,b in big if a<200), 'total us', round(sum(a for a,b in big if a<200)))
PY
