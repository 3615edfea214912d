#!/bin/bash
mkdir -p gpurun_out
timeout 600 bash profiles/ab_run.sh B E B E > gpurun_out/r2_ab4.txt 2>&1; cat gpurun_out/r2_ab4.txt
timeout 600 python profiles/hbm_kernels.py > gpurun_out/r2_hbm_kernels.jsonl 2> gpurun_out/r2_hbm.err; cat gpurun_out/r2_hbm_kernels.jsonl; tail -3 gpurun_out/r2_hbm.err
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_hbm_ncu.csv python profiles/hbm_kernels.py once > gpurun_out/r2_hbm_ncu.log 2>&1; tail -2 gpurun_out/r2_hbm_ncu.log; wc -l gpurun_out/r2_hbm_ncu.csv
