#!/bin/bash
# Extra evidence in one gpurun call: compute-sanitizer memcheck over the adaptive / puncturer / ragged-batch tests and
# a BLER sweep through the adaptive decoder next to the plain list decoder.  usage: tools/gpu_extra.sh [tag]
TAG=${1:-run}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 5 \
  python -m pytest tests/test_adaptive.py tests/test_puncturer.py -m gpu -q -x -p no:cacheprovider \
  -k "not 2048 and not 4096" > $OUT/${TAG}_memcheck.log 2>&1
echo "memcheck rc=$?"; tail -4 $OUT/${TAG}_memcheck.log
for MODE in "" "--adaptive"; do
  timeout 600 python -m antpolarcodes_b200.bler --N 1024 --K 512 --L 8 --crc 16 --construction 5G --ebn0 0:3:0.5 \
    --frames 400000 $MODE >> $OUT/${TAG}_bler_adaptive_vs_scl8.jsonl 2>> $OUT/${TAG}_bler.err
  echo "bler $MODE rc=$?"
done
cat $OUT/${TAG}_bler_adaptive_vs_scl8.jsonl | cut -c1-200
