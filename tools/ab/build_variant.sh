#!/bin/bash
# usage: build_variant.sh <name> [extra nvcc flags...]   -> tools/ab/<name>.so   (A/B experiments only)
set -e
name=$1; shift
d=/tmp/ab_$name; mkdir -p $d
for s in tacotron2_b200/csrc/*.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda -Xcompiler -fPIC -Xcompiler -O3 "$@" -c $s -o $d/$(basename $s).o &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o tools/ab/$name.so $d/*.o
