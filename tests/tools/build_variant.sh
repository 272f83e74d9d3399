#!/bin/bash
# Builds a tuning variant of libgpuhash.so for A/B runs: tests/tools/build_variant.sh <name> <extra nvcc -D flags...>
# -> build/libgpuhash_<name>.so (select with GPUHASH_LIB=...). build/ is git-ignored but travels to the GPU box.
set -e
cd "$(dirname "$0")/../.."
name=$1; shift
mkdir -p build/obj_$name
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-Wall,-Wno-unused-function"
src=downloader-go_b200/csrc
$NV "$@" -c -o build/obj_$name/gh_kernels.o $src/gh_kernels.cu
[ -f $src/gh_api.o ] || make -s -C $src
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o build/libgpuhash_$name.so build/obj_$name/gh_kernels.o $src/gh_api.o $src/gh_metainfo.o $src/gh_config.o $src/gh_copy.o -lpthread
echo build/libgpuhash_$name.so
