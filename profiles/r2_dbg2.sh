#!/bin/bash
# ncu --set full of one full-pass launch of the quadtree scoring kernel (dense full-class encode of lena 512x512)
timeout 600 ncu --set full --import-source on --clock-control none -k regex:qt_tcscore_kernel --launch-skip ${SKIP:-21} --launch-count 1 -f -o gpurun_out/r2_qtscore python -c "
import sys,os,numpy as np,time
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,True,False,4,8,2.0,1,1)
" > gpurun_out/ncu_qtscore.log 2>&1; tail -2 gpurun_out/ncu_qtscore.log
