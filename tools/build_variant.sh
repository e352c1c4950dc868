#!/bin/bash
# Tuning helper: rebuild only the 2-DOF instantiations with extra -D flags into
# aslr_to_b200/variants/<name>.so (the other objects come from the last full build).
#   tools/build_variant.sh gs32 -DASLR_BWD_GS=32 ;  ASLR_B200_LIB=aslr_to_b200/variants/gs32.so python bench.py
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p aslr_to_b200/variants build/obj
which=${ASLR_VARIANT_OF:-n2}   # n2 or n7: which instantiation file to rebuild
other=$([ "$which" = n2 ] && echo n7 || echo n2)
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include -I aslr_to_b200/csrc \
  "$@" -c aslr_to_b200/csrc/inst_$which.cu -o build/obj/inst_${which}_$name.o
nvcc -shared -o aslr_to_b200/variants/$name.so build/obj/api.o build/obj/inst_${which}_$name.o build/obj/inst_$other.o \
  -lcudart_static -lpthread -ldl -lrt
echo aslr_to_b200/variants/$name.so
