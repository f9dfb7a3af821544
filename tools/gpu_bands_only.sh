#!/bin/bash
# multi-GPU visit, band timing only: tools/bench_bands.py for the listed PSFs (env passes through, e.g. SR459_LAT_NB)
tag=${1:-r2x}; N=${2:-2}; shift; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
port=29530
for psf in "$@"; do
  port=$((port+1))
  timeout 300 $TR $port tools/bench_bands.py --psf $psf --reps 2 > gpurun_out/${tag}_bands_${psf}_n${N}.json 2> gpurun_out/${tag}_bands_${psf}_n${N}.err; echo "bands $psf rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/${tag}_bands_${psf}_n${N}.json').read().strip().splitlines()[-1])
print('$psf N=$N', round(d['us_per_iteration'],1), 'us/iter', [round(x,1) for x in d['per_rank_us_per_iteration']])
"
done
