#!/bin/bash
# build/libt2b200_<name>.so = the working tree's library with decoder_ar.cu compiled with extra -D flags (same-box A/B of kernel variants)
name=$1; shift
nvcc=/usr/local/cuda/bin/nvcc
$nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda -Xcompiler -fPIC -Xcompiler -O3 "$@" -c tacotron2_b200/csrc/decoder_ar.cu -o build/decoder_ar_$name.o || exit 1
objs=$(ls build/*.cu.o | grep -v decoder_ar.cu.o)
$nvcc -shared -gencode arch=compute_100a,code=sm_100a -o build/libt2b200_$name.so $objs build/decoder_ar_$name.o && echo built build/libt2b200_$name.so
