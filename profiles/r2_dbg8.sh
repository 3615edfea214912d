#!/bin/bash
# launch list of one default quadtree encode (lena 512x512)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_qt_default_launches.csv python -c "
import sys,os,numpy as np
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,False,False,4,8,2.0)
" > gpurun_out/ncu_qtd.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r2_qt_default_launches.csv')) if len(r)>10]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); ui=h.index('Metric Unit'); gi=h.index('Grid Size')
tot=0
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v = v/1e3 if u=='ns' else v
    tot+=v; print('%-46s %-18s %8.1f us'%(r[ki].split('(')[0][-46:], r[gi], v))
print('sum', tot)
PY
python - <<'PY'
import sys,time,numpy as np
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
h=fic_b200.default_handle(0)
for _ in range(3): fic_b200.qt_encode(img,7,5,1.0,False,False,4,8,2.0,handle=h)
t=[]
for _ in range(20):
    t0=time.perf_counter(); fic_b200.qt_encode(img,7,5,1.0,False,False,4,8,2.0,handle=h); t.append((time.perf_counter()-t0)*1e3)
print('wall ms median', np.median(t), 'min', min(t))
PY
