#!/bin/bash
# Builds the tuning variants that tools/gpu_ab.sh compares (variants/*.so are git-ignored but travel to the GPU box).
#   base.so            the default build (same-box reference for the A/B)
#   pair               explicit element kernels two elements per thread in packed fp32 (FFMA2), 128-thread CTAs
#   pair_nopf          packed pairs without the register prefetch of the next pair
#   persist_explicit   explicit tile kernel persistent, next tile's vertices by cp.async (scalar elements)
#   persist_pair       the same with packed pairs; the next tile's first planes requested before the gather
set -e
cd "$(dirname "$0")/../finite-element-method_b200/csrc"
mkdir -p ../../variants
make -j8 OUT=../../variants/base.so OBJDIR=build_base > /dev/null
make -j8 EXTRA=-DFEMB_PAIR_EXPLICIT=1 OUT=../../variants/pair.so OBJDIR=build_pair > /dev/null
make -j8 EXTRA="-DFEMB_PAIR_EXPLICIT=1 -DFEMB_PAIR_PREFETCH=0" OUT=../../variants/pair_nopf.so OBJDIR=build_nopf > /dev/null
make -j8 EXTRA="-DFEMB_PERSIST_EXPLICIT=1" OUT=../../variants/persist_explicit.so OBJDIR=build_pe > /dev/null
make -j8 EXTRA="-DFEMB_PERSIST_EXPLICIT=1 -DFEMB_PAIR_EXPLICIT=1" OUT=../../variants/persist_pair.so OBJDIR=build_pp > /dev/null
ls -la ../../variants
