#!/bin/bash
# list lengths of the quadtree exact pass, two library builds (debug hooks)
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
for v in "$@"; do
cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so
echo "== $v"
python -c "
import sys,os,numpy as np,time
sys.path.insert(0,'fractal-image-compression_b200')
import fic_b200
img=np.load('tests/golden/images.npz')['lena512']
fic_b200.qt_encode(img,7,5,1.0,True,False,4,8,2.0,1,1)
" 2>&1 | grep qtc | awk 'NR<=6 || /S=4 g0=(0|16|160|400) /'
done
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
