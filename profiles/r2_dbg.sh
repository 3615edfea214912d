#!/bin/bash
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
cp ab/libT8.so fractal-image-compression_b200/fic_b200/libfic_b200.so
FIC_DBG_LISTS=1 python -c "
import sys,os,numpy as np,time
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,True,False,4,8,2.0,1,1)
" 2>&1 | tail -12
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_qt_launches.csv python -c "
import sys,os,numpy as np,time
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,True,False,4,8,2.0,1,1)
" > /dev/null 2>&1
