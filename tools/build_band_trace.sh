#!/bin/bash
# builds tools/_bin/band_trace (with the phase stamps) and tools/_bin/band_notrace from tools/band_trace.cu
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/_bin
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -lineinfo -I wild-video-3d-reconstruction_b200/csrc -I include"
nvcc $F -DWV_BAND_TRACE tools/band_trace.cu -o tools/_bin/band_trace
nvcc $F -Xptxas -v tools/band_trace.cu -o tools/_bin/band_notrace 2>&1 | grep -A2 "ba_band" | grep "spill\|Used" || true
