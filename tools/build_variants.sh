#!/bin/bash
# Build tuning variants of the library: tools/build_variants.sh "64:6 32:12 ..." (sub bytes : overlap subs)
# Output: jpeggpu_b200/lib/variants/lib_s<S>_r<R>.so ; select with JPEGGPU_B200_LIB=<path>.
set -e
cd "$(dirname "$0")/../jpeggpu_b200/csrc"
mkdir -p ../lib/variants
for v in $1; do
  S=${v%%:*}; R=${v##*:}
  OUT=../lib/variants/lib_s${S}_r${R}.so
  nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-O3 \
    -I../../include -I. -DJB2_SUB_BYTES=$S -DJB2_OVERLAP_SUBS=$R -shared -Xlinker -Bsymbolic \
    -x cu kernels_destuff.cu kernels_huffman.cu kernels_dc.cu kernels_idct.cu decoder.cpp parser.cpp batch.cpp -lpthread -o $OUT &
done
wait
ls -la ../lib/variants
