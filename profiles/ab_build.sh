#!/bin/bash
# Build a variant of the library for same-box A/B timing: profiles/ab_build.sh NAME "-DFIC_EPI_ALL=0 ..." [TUNING=1]
# -> ab/libNAME.so (ab/ is git-ignored; it travels to the GPU box with the snapshot)
set -e
name=$1; extra=$2; shift 2 || true
root=$(cd "$(dirname "$0")/.." && pwd)
cs=$root/fractal-image-compression_b200/csrc
mkdir -p $root/ab
touch $cs/bf_screen.cu
make -C $cs -j8 -s EXTRA="$extra" "$@" >/dev/null
cp $root/fractal-image-compression_b200/fic_b200/libfic_b200.so $root/ab/lib$name.so
echo "built ab/lib$name.so with '$extra' $*"
