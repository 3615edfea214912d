#!/bin/bash
# launch list of one dense full-class quadtree encode (lena 512x512, ranges 4..8, tolerance 2.0)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_qt_dense_full_launches.csv python -c "
import sys,os,numpy as np,time
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,True,False,4,8,2.0,1,1)
" > gpurun_out/ncu_qtscore.log 2>&1; tail -2 gpurun_out/ncu_qtscore.log
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/r2_qt_dense_full_launches.csv')) if len(r)>10]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); ui=h.index('Metric Unit')
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v = v/1e3 if u=='ns' else (v if u in ('us','usecond') else v*1e3 if u=='ms' else v)
    k=r[ki].split('(')[0][-60:]; agg[k][0]+=1; agg[k][1]+=v
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:14]: print('%-62s %4d %10.1f us'%(k,n,t))
PY
