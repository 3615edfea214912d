#!/bin/bash
# N=8 bench under three settings, same box: default, NCCL limited to one CTA, no record gather (diagnostic).
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 --steps 50 --warmup 5 --no-cpu 2>> gpurun_out/n8_variants.err | tail -n 1; }
mkdir -p gpurun_out; : > gpurun_out/n8_variants.err
echo "default";        run 29601 > gpurun_out/n8_default.json
echo "one cta";        NCCL_MAX_CTAS=1 NCCL_MIN_CTAS=1 run 29602 > gpurun_out/n8_onecta.json
echo "no gather";      FIC_BENCH_NO_GATHER=1 run 29603 > gpurun_out/n8_nogather.json
python - <<'PY'
import json
for n in ("default", "onecta", "nogather"):
    try:
        d = json.loads(open(f"gpurun_out/n8_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["ms_per_step"], 4), [round(x, 3) for x in d["ms_per_step_by_rank"]])
    except Exception as e:
        print(n, "failed", e)
PY
