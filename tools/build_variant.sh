#!/bin/bash
# Builds a variant of the library with extra nvcc defines for A/B runs on the GPU box:
#   tools/build_variant.sh <tag> [-DJB_...=...]...   ->  build_variants/libjb_<tag>.so  (select with JBONSAI_B200_LIB)
set -e
TAG=$1; shift
cd "$(dirname "$0")/../jbonsai_b200/csrc"
mkdir -p ../../build_variants/obj_$TAG
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 -Xcompiler -fPIC"
for f in cuda/*.cu; do
  o=../../build_variants/obj_$TAG/$(basename $f .cu).o
  if [ "$f" = "cuda/mlsa_seg.cu" ] || [ ! -f $o ]; then $NV "$@" -c $f -o $o & fi
done
wait
make -s >/dev/null
$NV -shared -o ../../build_variants/libjb_$TAG.so ../../build_variants/obj_$TAG/*.o build/engine.o build/capi.o build/voice.o build/frontend.o -lpthread
echo built build_variants/libjb_$TAG.so
