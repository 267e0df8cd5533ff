#!/bin/bash
# usage: scripts/ptxas_summary.sh [extra nvcc flags]
# Rebuilds libvpt_cuda.so with -Xptxas -v and prints registers / spills of the wavefront kernels.
cd "$(dirname "$0")/../volumetric-path-tracer_b200"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v "$@" -shared -o libvpt_cuda.so csrc/vpt_cuda.cu 2>&1 | grep -A3 "Compiling entry" | grep -E "Compiling|registers|spill" | paste - - - | sed -E 's/ptxas info    ://g; s/Compiling entry function//; s/for .sm_100a.//' | awk '{ $1=$1; print }' | grep -E "k_wf|k_render_megaILi0|k_intersect" | sed -E 's/_ZN3vpt[0-9]+//; s/EvNS_6DScene[A-Za-z0-9_]*//' | sort
