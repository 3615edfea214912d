#!/bin/bash
# A/B timing of two builds of the library on the same box: ab/run.sh A B A B
for v in "$@"; do
  cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so
  echo "$v: $(python bench.py --steps 30 --warmup 4 --no-cpu 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['config']['parity'])")"
done
