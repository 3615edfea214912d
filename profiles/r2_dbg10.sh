#!/bin/bash
# launch list of one brute-force encode with 4x4 ranges (lena 512x512)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_bf4_launches.csv python -c "
import sys,os,numpy as np
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.bf_encode(img, range_size=4, domain_size=8)
" > gpurun_out/ncu_bf4.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r2_bf4_launches.csv')) if len(r)>10]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); ui=h.index('Metric Unit'); gi=h.index('Grid Size')
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v = v/1e3 if u=='ns' else v
    print('%-50s %-18s %10.1f us'%(r[ki].split('(')[0][-50:], r[gi], v))
PY
