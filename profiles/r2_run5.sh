#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest5.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest5.log
tail -25 gpurun_out/r2_pytest5.log
timeout 600 python profiles/run_configs.py > gpurun_out/r2_configs.jsonl 2> gpurun_out/r2_configs.err; cat gpurun_out/r2_configs.jsonl | cut -c1-600; tail -3 gpurun_out/r2_configs.err
timeout 600 bash profiles/ab_run.sh B S B S > gpurun_out/r2_ab5.txt 2>&1; cat gpurun_out/r2_ab5.txt
timeout 600 python profiles/hbm_kernels.py > gpurun_out/r2_hbm_kernels.jsonl 2> gpurun_out/r2_hbm.err; cat gpurun_out/r2_hbm_kernels.jsonl | cut -c1-400; tail -3 gpurun_out/r2_hbm.err
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_hbm_ncu.csv python profiles/hbm_kernels.py once > gpurun_out/r2_hbm_ncu.log 2>&1; tail -2 gpurun_out/r2_hbm_ncu.log
