#!/bin/bash
# profiles/build_variant.sh NAME [extra nvcc flags...] -> variants/NAME.so (a build of the library with extra -D flags, for A/B runs)
set -e
name=$1; shift
src=eags-slam_b200/csrc; tmp=$(mktemp -d)
F="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC"
for f in preprocess binning blend loss ssim pose c_api knn collective; do
  [ -f $src/$f.cu ] || continue
  extra=""; [ $f = preprocess ] && extra="-fmad=false"
  nvcc $F $extra "$@" -c $src/$f.cu -o $tmp/$f.o &
done
wait
mkdir -p variants
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o variants/$name.so $tmp/*.o
rm -rf $tmp; echo built variants/$name.so
