#!/bin/bash
# scaling visit (gpurun --gpus N): bench.py and the row-band IBP (config 4) at N GPUs, CUDA-graph loops
tag=${1:-r01x}; N=${2:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${N}.json 2> gpurun_out/${tag}_bench_n${N}.err; echo "bench N=$N rc=$?"
cut -c1-160 gpurun_out/${tag}_bench_n${N}.json
timeout 200 $TR tools/bench_bands.py --psf measured31 --iters 200 --check --graph 2>gpurun_out/${tag}_bands_m31_n${N}.err | grep -E "^\{" > gpurun_out/${tag}_bands_m31_n${N}.json; echo "bands m31 rc=$?"
cat gpurun_out/${tag}_bands_m31_n${N}.json
timeout 200 $TR tools/bench_bands.py --psf gaussian7 --iters 200 --check --graph 2>gpurun_out/${tag}_bands_g7_n${N}.err | grep -E "^\{" > gpurun_out/${tag}_bands_g7_n${N}.json; echo "bands g7 rc=$?"
cat gpurun_out/${tag}_bands_g7_n${N}.json
