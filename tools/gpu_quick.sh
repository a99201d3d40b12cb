#!/bin/bash
# One short gpurun call: a pytest selection + a few benches.  usage: tools/gpu_quick.sh TAG "pytest -k expr" "wl1 wl2" [extra bench args]
TAG=${1:-q}; KEXPR=${2:-}; WLS=${3:-scl8}; shift 3
OUT=gpurun_out; mkdir -p $OUT
if [ -n "$KEXPR" ]; then
  timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider --tb=short -x -k "$KEXPR" > $OUT/${TAG}_pytest.log 2>&1
  echo "pytest rc=$?"; tail -15 $OUT/${TAG}_pytest.log
fi
for WL in $WLS; do
  timeout 600 python bench.py --workload $WL --steps 5 --warmup 3 --also none "$@" > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err
  echo "bench $WL rc=$?"; tail -3 $OUT/${TAG}_bench_$WL.err
  python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench_$WL.json"))
    print("$WL value %.3f e2e %.3f frac %.4f cpu %s parity %s/%s gate %s" % (d["value"], d["e2e"]["value"], d["roofline"]["frac"], (d.get("cpu_baseline") or {}).get("value"), d["check"].get("parity_mismatches"), d["check"].get("parity_frames"), d.get("parity_gate")))
except Exception as e:
    print("no json", e)
PY
done
