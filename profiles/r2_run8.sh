#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest8.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest8.log
tail -4 gpurun_out/r2_pytest8.log
timeout 600 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null; tail -c 600 gpurun_out/r2_bench_reference.json
timeout 600 python profiles/run_configs.py > gpurun_out/r2_configs.jsonl 2> gpurun_out/r2_configs.err; wc -l gpurun_out/r2_configs.jsonl
timeout 600 python profiles/hbm_kernels.py > gpurun_out/r2_hbm_kernels.jsonl 2> gpurun_out/r2_hbm.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --sustain 0 --c4-side 0 --c5-images 0 > gpurun_out/r2_ncu_launches.log 2>&1; wc -l gpurun_out/r2_launches.csv
python profiles/one_encode.py > gpurun_out/one.log 2>&1 && timeout 600 ncu --set full --import-source on --clock-control none -k regex:bf_screen_kernel --launch-skip 3 --launch-count 3 -f -o gpurun_out/r2_screen python profiles/one_encode.py > gpurun_out/ncu_screen.log 2>&1; tail -1 gpurun_out/ncu_screen.log
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['kernel_ms_per_launch'], d['config']['parity'], d.get('c4',{}).get('seconds'), d.get('c5',{}).get('seconds'), d['sustained']['frac_of_sustained_peak'])
"
