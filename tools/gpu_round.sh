#!/bin/bash
# One gpurun call: GPU parity tests, benches of both headline workloads and (only when the tests are green)
# the ncu launch list + one full capture of the SCL kernel.  Everything lands in gpurun_out/.
# usage: tools/gpu_round.sh [tag]   (run from the repo root on the GPU box)
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider --tb=short > $OUT/${TAG}_pytest.log 2>&1
RC=$?
tail -5 $OUT/${TAG}_pytest.log
echo "pytest rc=$RC"
for WL in ${BENCH_WLS:-scl8 fastssc fastssc_f32}; do
  timeout 300 python bench.py --workload $WL --steps 5 --warmup 3 > $OUT/${TAG}_bench_$WL.json 2> $OUT/${TAG}_bench_$WL.err
  echo "bench $WL rc=$?"; cat $OUT/${TAG}_bench_$WL.json | cut -c1-600
done
# optional A/B runs: VARIANTS="scl8:POLAR_CUDA_LOCKSTEP=0 fastssc:POLAR_CUDA_X=1"
for V in ${VARIANTS:-}; do
  WL=${V%%:*}; KV=${V#*:}
  env $KV timeout 300 python bench.py --workload $WL --steps 5 --warmup 3 > $OUT/${TAG}_bench_${WL}_${KV}.json 2> $OUT/${TAG}_bench_${WL}_${KV}.err
  echo "variant $WL $KV rc=$?"; python -c "import json,sys; d=json.load(open('$OUT/${TAG}_bench_${WL}_${KV}.json')); print(d['value'], d['e2e']['value'], d['config']['warps_per_block'], d['config']['blocks_per_sm'])"
done
if [ $RC -eq 0 ] && [ "${NCU:-1}" = "1" ]; then
  for WL in ${NCU_WLS:-scl8 fastssc}; do
    if [ $WL = scl8 ]; then KRN=scl_i8_kernel; FR=131072; elif [ $WL = scl8_f32 ]; then KRN=scl_f32_kernel; FR=65536; elif [ $WL = encode ]; then KRN=encode_thread_kernel; FR=4194304; elif [ $WL = fastssc_f32 ]; then KRN=fastssc_f32_kernel; FR=262144; else KRN=fastssc_i8_kernel; FR=262144; fi
    CMD="python bench.py --workload $WL --steps 2 --warmup 1 --frames $FR --e2e-frames 8192"
    $CMD > $OUT/${TAG}_ncu_plain_$WL.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches_$WL.csv $CMD > $OUT/${TAG}_ncu_list_$WL.log 2>&1
    ncu --set full --clock-control none --import-source on -k regex:$KRN -s 1 -c 1 -o $OUT/${TAG}_prof_$WL -f $CMD > $OUT/${TAG}_ncu_full_$WL.log 2>&1
    echo "ncu $WL rc=$?"
  done
fi
