#!/bin/bash
# tools/variant_rev.sh NAME REV [nvcc flags]: A/B build of the CUDA layer as of git revision REV (see variant.sh)
set -e
cd "$(dirname "$0")/.."
name=$1; rev=$2; shift 2
tmp=$(mktemp -d)
mkdir -p $tmp/csrc $tmp/include desamba-so_b200/lib_$name
for f in dsb_core.cuh dsb_layout.cuh dsb_cuda.cu; do git show $rev:desamba-so_b200/csrc/$f > $tmp/csrc/$f; done
git show $rev:include/dsb_cuda.h > $tmp/include/dsb_cuda.h
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda \
    -Xcompiler -fPIC -I $tmp/include -I $tmp/csrc "$@" -shared -o desamba-so_b200/lib_$name/libdsb_cuda.so $tmp/csrc/dsb_cuda.cu
cp desamba-so_b200/lib/libdesamba.so desamba-so_b200/lib_$name/
rm -rf $tmp
echo "built lib_$name from $rev"
