#!/bin/bash
# e2e rate of the headline against the chunk sizes of the HOST pipeline.  usage: tools/exp_e2e_chunks.sh "K=V K2=V2" ...
for KV in "$@"; do
 env $KV python bench.py --workload ${WL:-scl8} --also none --no-cpu --parity 4096 --steps 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$KV value %.3f e2e %.3f' % (d['value'], d['e2e']['value']))"
done
