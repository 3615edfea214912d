#!/bin/bash
cp fractal-image-compression_b200/fic_b200/libfic_b200.so ab/.shipped.so
for v in S13 HP4 HP6 HP8 S13; do cp ab/lib$v.so fractal-image-compression_b200/fic_b200/libfic_b200.so; echo "$v $(python profiles/tree_timing.py 2>&1 | tail -1 | cut -c1-45)"; done
cp ab/.shipped.so fractal-image-compression_b200/fic_b200/libfic_b200.so
