#!/bin/bash
# round 2, first GPU visit: parity suite, bench, A/B of the screen-kernel variants, microbenchmarks
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/r2_gpu.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest1.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest1.log
tail -5 gpurun_out/r2_pytest1.log
timeout 600 python bench.py > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/r2_bench1.json; tail -5 gpurun_out/r2_bench1.err
timeout 900 bash profiles/ab_run.sh R1 B C C8 R1 B C C8 > gpurun_out/r2_ab1.txt 2>&1; cat gpurun_out/r2_ab1.txt
timeout 600 bash profiles/ab_env.sh "FIC_SCREEN_DEBUG=0" "FIC_SCREEN_DEBUG=1" "FIC_SCREEN_DEBUG=2" "FIC_SCREEN_DEBUG=3" > gpurun_out/r2_abenv1.txt 2>&1; cat gpurun_out/r2_abenv1.txt
(cd profiles/microbench && timeout 120 ./tmem_ld_bw > ../../gpurun_out/r2_tmem_ld_bw.txt 2>&1; timeout 120 ./alu_rates > ../../gpurun_out/r2_alu_rates.txt 2>&1)
tail -20 gpurun_out/r2_tmem_ld_bw.txt gpurun_out/r2_alu_rates.txt
