#!/bin/bash
# ncu --set full of one large HV search launch (lena 512x512)
timeout 600 ncu --set full --import-source on --clock-control none -k regex:hv_search --launch-skip ${SKIP:-14} --launch-count 1 -f -o gpurun_out/r2_hvsearch python profiles/hv_once.py 1 > gpurun_out/ncu_hv2.log 2>&1; tail -1 gpurun_out/ncu_hv2.log
