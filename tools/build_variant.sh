#!/bin/bash
# build/libt2b200_<name>.so = the working tree's library with ONE source (SRC, default decoder_ar.cu) compiled with extra -D flags
# (same-box A/B of kernel variants).  usage: [SRC=bwd_decoder.cu] tools/build_variant.sh name -DFLAG...
name=$1; shift
src=${SRC:-decoder_ar.cu}
stem=${src%.cu}
nvcc=/usr/local/cuda/bin/nvcc
$nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda -Xcompiler -fPIC -Xcompiler -O3 "$@" -c tacotron2_b200/csrc/$src -o build/${stem}_$name.o || exit 1
objs=$(ls build/*.cu.o | grep -v "/$src.o")
$nvcc -shared -gencode arch=compute_100a,code=sm_100a -o build/libt2b200_$name.so $objs build/${stem}_$name.o && echo built build/libt2b200_$name.so
