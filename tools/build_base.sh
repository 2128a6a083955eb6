#!/bin/bash
# Build the library of another revision (default HEAD) as build/libt2b200_base.so for same-box A/B runs (T2B200_LIB=build/libt2b200_base.so).
rev=${1:-HEAD}
rm -rf /tmp/t2base && mkdir -p /tmp/t2base && git archive $rev tacotron2_b200/csrc include | tar -x -C /tmp/t2base
objs=()
for s in /tmp/t2base/tacotron2_b200/csrc/*.cu; do
  o=/tmp/t2base/$(basename $s).o; objs+=($o)
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda -Xcompiler -fPIC -Xcompiler -O3 -c $s -o $o &
done
wait
mkdir -p build
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o build/libt2b200_base.so "${objs[@]}" && ls -la build/libt2b200_base.so
