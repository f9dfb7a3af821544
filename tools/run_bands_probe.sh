#!/bin/bash
# 7x7 row bands: where does the exchange cost go?  SR459_BAND_PROBE bits: 1 no arrival waits, 2 no peer stores, 4 no fence / arrival publish
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
port=29550
for probe in 0 1 2 4 3 6 7; do
  port=$((port+1))
  SR459_BAND_PROBE=$probe timeout 200 $T $port tools/bench_bands.py --psf gaussian7 --reps 2 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('probe=$probe', round(d['us_per_iteration'],2), 'us/iter', [round(x,1) for x in d['per_rank_us_per_iteration']])"
done
port=$((port+1))
SR459_BAND_NOPUSH=1 timeout 200 $T $port tools/bench_bands.py --psf gaussian7 --reps 2 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('nopush', round(d['us_per_iteration'],2), 'us/iter', [round(x,1) for x in d['per_rank_us_per_iteration']])"
