#!/bin/bash
timeout 600 ncu --set full --import-source on --clock-control none -k regex:bf_screen_kernel --launch-skip 1 --launch-count 1 -f -o gpurun_out/r2_bf4 python -c "
import sys,os,numpy as np
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.bf_encode(img, range_size=4, domain_size=8)
" > gpurun_out/ncu_bf4b.log 2>&1; tail -1 gpurun_out/ncu_bf4b.log
