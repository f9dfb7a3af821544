#!/bin/bash
# multi-GPU visit (gpurun --gpus N): bench.py under torchrun, row-band IBP (config 4) with parity check
tag=${1:-r01x}; N=${2:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${N}.json 2> gpurun_out/${tag}_bench_n${N}.err; echo "bench N=$N rc=$?"
cat gpurun_out/${tag}_bench_n${N}.json; tail -3 gpurun_out/${tag}_bench_n${N}.err
timeout 600 $TR bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/${tag}_bench_ref_n${N}.json 2> gpurun_out/${tag}_bench_ref_n${N}.err; echo "bench ref N=$N rc=$?"
timeout 600 $TR tools/bench_bands.py --psf gaussian7 --iters 200 --check > gpurun_out/${tag}_bands_g7_n${N}.json 2> gpurun_out/${tag}_bands_g7_n${N}.err; echo "bands g7 rc=$?"
cat gpurun_out/${tag}_bands_g7_n${N}.json; tail -3 gpurun_out/${tag}_bands_g7_n${N}.err
timeout 900 $TR tools/bench_bands.py --psf measured31 --iters 200 --check > gpurun_out/${tag}_bands_m31_n${N}.json 2> gpurun_out/${tag}_bands_m31_n${N}.err; echo "bands m31 rc=$?"
cat gpurun_out/${tag}_bands_m31_n${N}.json; tail -3 gpurun_out/${tag}_bands_m31_n${N}.err
timeout 600 python tools/bench_bands.py --psf gaussian7 --iters 200 > gpurun_out/${tag}_bands_g7_n1.json 2>&1; cat gpurun_out/${tag}_bands_g7_n1.json
timeout 900 python tools/bench_bands.py --psf measured31 --iters 200 > gpurun_out/${tag}_bands_m31_n1.json 2>&1; cat gpurun_out/${tag}_bands_m31_n1.json
