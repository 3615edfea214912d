#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest2.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest2.log
tail -5 gpurun_out/r2_pytest2.log
python profiles/one_encode.py > gpurun_out/one.log 2>&1 && \
timeout 600 ncu --set full --import-source on --clock-control none -k regex:bf_screen_kernel --launch-skip 3 --launch-count 3 -f -o gpurun_out/r2_screenB python profiles/one_encode.py > gpurun_out/ncuB.log 2>&1
tail -3 gpurun_out/ncuB.log
