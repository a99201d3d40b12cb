#!/bin/bash
# A/B runs of one workload under environment variants.  usage: WL=scl8 tools/exp_var.sh "TAG1:K=V,K2=V2" "TAG2:..." ...
OUT=gpurun_out; mkdir -p $OUT
for V in "$@"; do
  TAG=${V%%:*}; KV=${V#*:}; KV=${KV//,/ }
  env $KV timeout 300 python bench.py --workload ${WL:-scl8} --steps 5 --warmup 3 --also none --no-cpu --parity ${PAR:-4096} > $OUT/var_$TAG.json 2> $OUT/var_$TAG.err
  python - <<PY
import json
try:
    d=json.load(open("$OUT/var_$TAG.json"))
    print("$TAG [$KV] value %.3f warps %d x %d smem/warp %d parity %s" % (d["value"], d["config"]["warps_per_block"], d["config"]["blocks_per_sm"], d["config"]["smem_per_warp"], d["check"].get("parity_mismatches")))
except Exception as e:
    print("$TAG no json", e)
PY
done
