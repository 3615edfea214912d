#!/bin/bash
# same-box timing of one library build under different environment knobs: profiles/ab_env.sh "FIC_ROUND_GROWTH=8" "FIC_ROUND_GROWTH=16" ...
for e in "$@"; do
  echo "$e: $(env $e python bench.py --steps 30 --warmup 4 --no-cpu 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['launches_per_step'], d['screen']['survivors'], d['config']['parity'])")"
done
