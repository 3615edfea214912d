#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tree_gpu.py -x -q > gpurun_out/r2_pytest6.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2_pytest6.log
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
for v in E2 P1 E2 P1; do cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so; echo "$v $(python profiles/tree_timing.py 2>&1 | tail -1)"; done | tee gpurun_out/r2_tree_ab.txt
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
