#!/bin/bash
# ncu --set full of one launch of the MODE 6 GEMM (quadtree exact pass, 4x4 ranges) in a dense full-class encode of lena 512x512
timeout 600 ncu --set full --import-source on --clock-control none -k regex:bf_screen_kernel --launch-skip ${SKIP:-3} --launch-count 1 -f -o gpurun_out/r2_qtgemm python -c "
import sys,os,numpy as np,time
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,True,False,4,8,2.0,1,1)
" > gpurun_out/ncu_qtgemm.log 2>&1; tail -2 gpurun_out/ncu_qtgemm.log
