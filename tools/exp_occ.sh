#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
run() { # tag env...
  TAG=$1; shift
  env "$@" timeout 300 python bench.py --workload ${WL:-scl8_n512} --steps 5 --warmup 3 --also none --no-cpu --parity 4096 > $OUT/exp_$TAG.json 2> $OUT/exp_$TAG.err
  python - <<PY
import json
try:
    d=json.load(open("$OUT/exp_$TAG.json"))
    print("$TAG value %.3f warps %d x blocks %d parity %s" % (d["value"], d["config"]["warps_per_block"], d["config"]["blocks_per_sm"], d["check"].get("parity_mismatches")))
except Exception as e:
    print("$TAG no json", e)
PY
}
run A_default X=1
run A_w12_b1 POLAR_CUDA_WARPS_PER_BLOCK=12 POLAR_CUDA_MAX_BLOCKS_PER_SM=1
run A_w16_b1 POLAR_CUDA_WARPS_PER_BLOCK=16 POLAR_CUDA_MAX_BLOCKS_PER_SM=1
M=$PWD/antpolarcodes_b200/libpolar_cuda_mb2.so
run B_w12_b1 POLAR_CUDA_LIB=$M POLAR_CUDA_WARPS_PER_BLOCK=12 POLAR_CUDA_MAX_BLOCKS_PER_SM=1
run B_w12_b2 POLAR_CUDA_LIB=$M POLAR_CUDA_WARPS_PER_BLOCK=12 POLAR_CUDA_MAX_BLOCKS_PER_SM=2
run B_w16_b1 POLAR_CUDA_LIB=$M POLAR_CUDA_WARPS_PER_BLOCK=16 POLAR_CUDA_MAX_BLOCKS_PER_SM=1
run B_w8_b3 POLAR_CUDA_LIB=$M POLAR_CUDA_WARPS_PER_BLOCK=8 POLAR_CUDA_MAX_BLOCKS_PER_SM=3
run B_w10_b3 POLAR_CUDA_LIB=$M POLAR_CUDA_WARPS_PER_BLOCK=10
