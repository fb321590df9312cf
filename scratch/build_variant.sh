#!/bin/bash
# usage: build_variant.sh name [-DFLAG ...]   -> scratch/variants/lib_<name>.so
name=$1; shift
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared --use_fast_math -Xptxas -v "$@" -o scratch/variants/lib_$name.so smsgate_b200/csrc/smsgate_b200.cu 2>&1 | grep -A3 "Compiling entry function '_Z11k_scan_keys" | grep -E "spill|Used"
