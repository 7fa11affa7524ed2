#!/bin/bash
# Build a variant of the CUDA library with extra nvcc flags (e.g. -DTB_RES_GQ=16) into build_variants/libtb_<name>.so;
# run it with TB_LIB_PATH=build_variants/libtb_<name>.so python bench.py ...   usage: tools/build_variant.sh <name> <flags...>
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$root/build_variants"
cd "$root/textblaster_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" -o "$root/build_variants/libtb_$name.so" tb_api.cu tb_kernels.cu tb_noalloc.cu
echo "built build_variants/libtb_$name.so"
