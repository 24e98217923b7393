#!/bin/bash
# Build a tuning variant of libdftd4.so with extra nvcc -D flags: tools/build_variant.sh <tag> -DA2_KC=96 ...
# Output: dftd4_b200/lib/var/libdftd4_<tag>.so (select it with DFTD4_B200_LIB=...; git-ignored, ships with gpurun).
set -e
TAG=$1; shift
cd "$(dirname "$0")/.."
mkdir -p dftd4_b200/lib/var
OBJ=dftd4_b200/lib/var/engine_$TAG.o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC "$@" -c dftd4_b200/csrc/d4_engine.cu -o $OBJ
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o dftd4_b200/lib/var/libdftd4_$TAG.so $OBJ dftd4_b200/lib/obj/d4_host.cpp.o dftd4_b200/lib/obj/d4_api.cpp.o -ldl
rm -f $OBJ
echo built dftd4_b200/lib/var/libdftd4_$TAG.so
