#!/bin/bash
# Weak-scaling run of one workload on an N-GPU box (gpurun --gpus 8 -- 'bash tools/scale_run.sh scl8 r1'):
# the same torchrun launch the driver uses, N = 1, 2, 4, 8, one JSON line each into gpurun_out/<tag>_scale<N>_<wl>.json
WL=${1:-scl8}; TAG=${2:-run}; OUT=gpurun_out; mkdir -p $OUT
NG=$(nvidia-smi -L | wc -l)
for N in 1 2 4 8; do
  [ $N -gt $NG ] && break
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --workload $WL --steps 5 --warmup 3 > $OUT/${TAG}_scale${N}_$WL.json 2> $OUT/${TAG}_scale${N}_$WL.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $WL --steps 5 --warmup 3 > $OUT/${TAG}_scale${N}_$WL.json 2> $OUT/${TAG}_scale${N}_$WL.err
  fi
  echo "N=$N rc=$?"; python -c "import json;d=json.loads(open('$OUT/${TAG}_scale${N}_$WL.json').read().strip().splitlines()[-1]);print(d['n_gpus'],round(d['value'],2),round(d['e2e']['value'],2),d['check'])"
done
