#!/bin/bash
# two GPUs: the tests that need a second device, and the bench under torchrun
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -k "two or gpus or async or multi" > gpurun_out/r2_pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "rc=$?"; tail -2 gpurun_out/r2_bench_n2.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n2.json').read().strip().splitlines()[-1]); print(d['n_gpus'], '%.4g'%d['value'], '%.4f'%d['ms_per_step'], '%.4g'%d['e2e']['value'], ['%.3f'%x for x in d['ms_per_step_by_rank']], 'c4', d.get('c4',{}).get('seconds'), d.get('c4',{}).get('equals_single_gpu_slice'), 'c5', d.get('c5',{}).get('seconds'), d['config'].get('parity'))
"
