#!/bin/bash
# Host parser under sanitizers (CPU only, no GPU needed):
#  1. ASan + UBSan over a corpus of ~1300 mutated files (exact-size heap buffers, so any over-read
#     of the input is reported): eager walk vs deferred walk vs deferred walk with helper threads
#     must agree in status and directories (tests/host_emul/emul.cpp: emul_walk_compare).
#  2. TSan over 200 repetitions of the parallel walk on the 12 MP workload.
# Last run (round 1): files=1277 diffs=0, no sanitizer report; TSan: done bad=0, no report.
set -e
cd "$(dirname "$0")/.."
T=${TMPDIR:-/tmp}/jb2_sanitize
mkdir -p $T
python tools/sanitize/make_corpus.py $T/corpus
INC="-Ijpeggpu_b200/csrc -Iinclude -I/usr/local/cuda/include"
SRC="tests/host_emul/emul.cpp jpeggpu_b200/csrc/parser.cpp -lpthread"
g++ -O1 -g -std=c++17 -w -fsanitize=address,undefined -fno-sanitize-recover=undefined $INC \
    tools/sanitize/walk_corpus_main.cpp $SRC -o $T/asan_walk
ls $T/corpus/*.jpg | xargs $T/asan_walk | tail -3
g++ -O1 -g -std=c++17 -w -fsanitize=thread $INC tools/sanitize/walk_repeat_main.cpp $SRC -o $T/tsan_walk
python -c "import sys; sys.path[:0]=['tools']; import gen_inputs; print(gen_inputs.ensure_synth(4032, 3024, 6510))" > $T/w.txt
$T/tsan_walk "$(tail -1 $T/w.txt)" | tail -2
