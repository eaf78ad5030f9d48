#!/bin/bash
# Builds the tuning variants that tools/gpu_ab.sh compares (variants/*.so are git-ignored but travel to the GPU box).
#   base.so            the default build (same-box reference for the A/B)
#   persist_explicit   explicit tile kernel persistent, next tile's vertices by cp.async
set -e
cd "$(dirname "$0")/../finite-element-method_b200/csrc"
mkdir -p ../../variants
make -j8 OUT=../../variants/base.so OBJDIR=build_base > /dev/null
make -j8 EXTRA=-DFEMB_PERSIST_EXPLICIT=1 OUT=../../variants/persist_explicit.so OBJDIR=build_pe > /dev/null
ls -la ../../variants
