#!/bin/bash
mkdir -p gpurun_out
for v in V3 V4 V5; do
cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so
echo "== $v"; timeout 120 python profiles/one_encode.py 1 2>&1 | tail -1
done
