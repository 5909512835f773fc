#!/bin/bash
# Diagnostic build with kernel timeline stamps (-DJB2_TIMELINE): jpeggpu_b200/lib/diag/lib_timeline.so
set -e
cd "$(dirname "$0")/../jpeggpu_b200/csrc"
mkdir -p ../lib/diag
nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-O3 \
  -I../../include -I. -DJB2_TIMELINE $EXTRA_DEFS -shared -Xlinker -Bsymbolic \
  -x cu kernels_destuff.cu kernels_huffman.cu kernels_dc.cu kernels_idct.cu decoder.cpp parser.cpp batch.cpp -lpthread \
  -o ../lib/diag/lib_timeline.so
ls -la ../lib/diag
