#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo "rc=$?"; tail -3 gpurun_out/r2_bench_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --c5-images 0 > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err; echo "rc=$?"
timeout 300 python bench.py --no-cpu --c5-images 0 --sustain 0 > gpurun_out/r2_bench_n1_box8.json 2>/dev/null
python -c "
import json
for f in ('gpurun_out/r2_bench_n1_box8.json','gpurun_out/r2_bench_n4.json','gpurun_out/r2_bench_n8.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], '%.4g'%d['value'], '%.4f'%d['ms_per_step'], '%.4g'%d['e2e']['value'], ['%.3f'%x for x in d['ms_per_step_by_rank']], '%.4f'%d['roofline']['kernel_ms_per_launch'], 'c4', d.get('c4',{}).get('seconds'), d.get('c4',{}).get('equals_single_gpu_slice'), 'c5', d.get('c5',{}).get('seconds'))
"
