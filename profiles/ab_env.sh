#!/bin/bash
# same-box timing of one TUNING build (ab/libT.so, `ab_build.sh T "" TUNING=1`) under different environment knobs:
#   profiles/ab_env.sh "FIC_SCREEN_DEBUG=1" "FIC_SCREEN_DEBUG=2" ...
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
cp ab/libT.so fractal-image-compression_b200/fic_b200/libfic_b200.so
for e in "$@"; do
  echo "$e: $(env $e python bench.py --steps 30 --warmup 4 --no-cpu --sustain 0 --c4-side 0 --c5-images 0 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['launches_per_step'], d['screen']['survivors'], d['config']['parity'])")"
done
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
