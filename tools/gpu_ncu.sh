#!/bin/bash
# One gpurun call: plain run, then the ncu launch list and one full capture of a kernel.
# usage: tools/gpu_ncu.sh TAG WORKLOAD KERNEL_REGEX FRAMES
TAG=$1; WL=$2; KRN=$3; FR=${4:-131072}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --workload $WL --steps 2 --warmup 1 --frames $FR --e2e-frames 8192 --also none --no-cpu --parity 0"
$CMD > $OUT/${TAG}_ncu_plain_$WL.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches_$WL.csv $CMD > $OUT/${TAG}_ncu_list_$WL.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:$KRN -s 1 -c 1 -o $OUT/${TAG}_prof_$WL -f $CMD > $OUT/${TAG}_ncu_full_$WL.log 2>&1
echo "ncu $WL rc=$?"; ls -la $OUT/${TAG}_prof_$WL.ncu-rep
