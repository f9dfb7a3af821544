#!/bin/bash
# multi-GPU visit: bench.py under torchrun only (its `bands` record = config 4 as row bands, both PSFs)
tag=${1:-r2x}; N=${2:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
timeout 900 $TR 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${N}.json 2> gpurun_out/${tag}_bench_n${N}.err; echo "bench N=$N rc=$?"
tail -c 400 gpurun_out/${tag}_bench_n${N}.json; tail -3 gpurun_out/${tag}_bench_n${N}.err
