#!/bin/bash
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 600 python -m pytest tests/test_async_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "rc=$?"; tail -c 2500 gpurun_out/r2_bench_n2.json; tail -5 gpurun_out/r2_bench_n2.err
timeout 600 python bench.py --no-cpu --c4-side 0 --c5-images 0 > gpurun_out/r2_bench_n1_same_box.json 2>/dev/null; python -c "
import json
for f in ('gpurun_out/r2_bench_n1_same_box.json','gpurun_out/r2_bench_n2.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['ms_per_step_by_rank'], d['roofline']['kernel_ms_per_launch'])
"
