#!/bin/bash
# Builds libvpt_cuda.so with extra defines into scripts/_variants/<name>/ (git-ignored) for A/B timing on
# the GPU box: VPT_LIB_DIR=scripts/_variants/<name> python scripts/profile_step.py
# usage: scripts/build_variant.sh <name> "<nvcc defines>"
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
PKG=$ROOT/volumetric-path-tracer_b200
OUT=$ROOT/scripts/_variants/$1
mkdir -p "$OUT"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false \
  -Xcompiler -fPIC,-ffp-contract=off $2 -shared -o "$OUT/libvpt_cuda.so" "$PKG/csrc/vpt_cuda.cu"
cp "$PKG/libvpt_host.so" "$OUT/"
