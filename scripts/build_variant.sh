#!/bin/bash
# scripts/build_variant.sh <name> <extra nvcc flags...>  -> build/variants/<name>.so
name=$1; shift
mkdir -p build/variants
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC -shared "$@" \
  -o build/variants/$name.so jpeg_decoder_b200/csrc/b200jpeg.cu jpeg_decoder_b200/csrc/host_parse.cpp
