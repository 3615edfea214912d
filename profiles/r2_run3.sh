#!/bin/bash
mkdir -p gpurun_out
cp ab/libD3.so fractal-image-compression_b200/fic_b200/libfic_b200.so
timeout 600 python -m pytest tests/test_screen_gpu.py tests/test_async_gpu.py -x -q -m gpu > gpurun_out/r2_pytest3.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest3.log
timeout 900 bash profiles/ab_run.sh B D3 V1 B D3 V1 > gpurun_out/r2_ab3.txt 2>&1; cat gpurun_out/r2_ab3.txt
