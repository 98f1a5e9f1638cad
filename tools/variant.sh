#!/bin/bash
# tools/variant.sh NAME [nvcc -D flags...]: an A/B build of the CUDA layer in desamba-so_b200/lib_NAME/
# (git-ignored; select it at run time with DSB_LIB_DIR=desamba-so_b200/lib_NAME).  Timings of two variants are
# only comparable inside ONE gpurun call: boxes differ by 10 % and more.
set -e
cd "$(dirname "$0")/../desamba-so_b200"
name=$1; shift
mkdir -p lib_$name
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda \
    -Xcompiler -fPIC -I ../include -I csrc "$@" -shared -o lib_$name/libdsb_cuda.so csrc/dsb_cuda.cu obj/dsb_build.o -lz
cp lib/libdesamba.so lib_$name/
echo "built lib_$name ($*)"
