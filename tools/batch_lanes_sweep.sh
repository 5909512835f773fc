#!/bin/bash
# Batch throughput on one GPU against the number of lanes (host thread + decoder + stream each),
# for 2 MP images (BASELINE config 5) and for tiny 64x64 images, where the device work is
# negligible and the rate is what the host side (parse + 2 copies + event + 5 launches per image,
# driver calls of one process contending for the context) can enqueue.
for wh in "1920 1080" "64 64"; do
  set -- $wh
  for l in 1 2 4 8 16; do
    python tools/batch_decode.py --images 4096 --distinct 8 --width $1 --height $2 --lanes $l 2>/dev/null | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('${1}x${2} lanes', $l, 'images/s', d['value'], 'us/image', round(1e6/d['value'],1))"
  done
done
echo host cores: $(nproc)
