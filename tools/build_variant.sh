#!/bin/bash
# build a variant of the library with extra -D flags for ONE source (experiments): tools/build_variant.sh name src.cu -DX=1 ...
set -e
NAME=$1; SRC=$2; shift 2
PKG=1d-burgers-equation-roms_b200
mkdir -p tools/bin
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC "$@" -c -o tools/bin/$NAME.o $PKG/csrc/$SRC
OBJS=$(ls $PKG/build/*.o | grep -v "/${SRC%.cu}.o")
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o tools/bin/lib_$NAME.so $OBJS tools/bin/$NAME.o
rm tools/bin/$NAME.o
echo tools/bin/lib_$NAME.so
