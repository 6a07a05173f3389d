#!/bin/bash
# profiles/build_variant.sh NAME [extra nvcc flags...] -> variants/NAME.so
# A build of the library with extra -D flags, for A/B runs (EAGS_RASTER_LIB=variants/NAME.so); same
# Makefile, own object directory.
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p $root/variants
make -s -C $root/eags-slam_b200/csrc -j8 OBJDIR=$root/variants/obj_$name OUT=$root/variants/$name.so EXTRA="$*"
echo built variants/$name.so
