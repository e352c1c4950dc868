#!/bin/bash
# Tuning helper: rebuild only one link-DOF instantiation with extra -D flags into
# aslr_to_b200/variants/<name>.so (the other objects come from the last full build).
#   tools/build_variant.sh gs32 -DASLR_BWD_GS=32 ;  ASLR_B200_LIB=aslr_to_b200/variants/gs32.so python bench.py
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p aslr_to_b200/variants build/obj
which=${ASLR_VARIANT_OF:-n2}   # n1 .. n7: which instantiation file to rebuild
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include -I aslr_to_b200/csrc \
  "$@" -c aslr_to_b200/csrc/inst_$which.cu -o build/obj/inst_${which}_$name.o
others=$(for n in 1 2 3 4 5 6 7; do [ "n$n" = "$which" ] || echo build/obj/inst_n$n.o; done)
nvcc -shared -o aslr_to_b200/variants/$name.so build/obj/api.o build/obj/inst_${which}_$name.o $others \
  -lcudart_static -lpthread -ldl -lrt
echo aslr_to_b200/variants/$name.so
