#!/bin/bash
# launch list of a short sweep point: share of the decoder kernel in the simulator chain
OUT=gpurun_out; mkdir -p $OUT
CMD="python -m antpolarcodes_b200.bler --N ${N:-1024} --configs ${CFG:-8:16} --ebn0 3.5 --max-frames 1048576 --round-frames 1048576 --no-resume"
$CMD > $OUT/${1}_sim_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${1}_sim_launches.csv $CMD > $OUT/${1}_sim_ncu.log 2>&1
python tools/launch_shares.py $OUT/${1}_sim_launches.csv | head -20
