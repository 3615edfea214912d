#!/bin/bash
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
for v in "$@"; do cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so; echo "$v $(python profiles/rgb_qt_timing.py 2>&1 | tail -1)"; done
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
