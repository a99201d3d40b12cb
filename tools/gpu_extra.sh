#!/bin/bash
# Extra evidence in one gpurun call: a BLER sweep through the adaptive decoder next to the plain list decoder.  usage: tools/gpu_extra.sh [tag]
TAG=${1:-run}; OUT=gpurun_out; mkdir -p $OUT
# (compute-sanitizer is closed on this pool: bounds are covered by the ragged / tiled / all-fail parity cases instead)
for MODE in "" "--adaptive"; do
  timeout 600 python -m antpolarcodes_b200.bler --N 1024 --K 512 --L 8 --crc 16 --construction 5G --ebn0 0:3:0.5 \
    --frames 400000 $MODE >> $OUT/${TAG}_bler_adaptive_vs_scl8.jsonl 2>> $OUT/${TAG}_bler.err
  echo "bler $MODE rc=$?"
done
cat $OUT/${TAG}_bler_adaptive_vs_scl8.jsonl | cut -c1-200
