#!/bin/bash
# A/B timing of library builds on the same box: profiles/ab_run.sh A B A B   (ab/libA.so, ab/libB.so from ab_build.sh)
# The shipped library is put back afterwards.
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
for v in "$@"; do
  cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so
  echo "$v: $(python bench.py --steps 30 --warmup 4 --no-cpu --sustain 0 --c4-side 0 --c5-images 0 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['screen']['survivors'], d['screen']['ranges_rescanned'], d['config']['parity'])")"
done
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
