#!/bin/bash
# multi-GPU visit (gpurun --gpus N): bench.py under torchrun (its `bands` record = config 4 as row bands) + band timing tool
tag=${1:-r2x}; N=${2:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
timeout 600 $TR 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${N}.json 2> gpurun_out/${tag}_bench_n${N}.err; echo "bench N=$N rc=$?"
tail -c 300 gpurun_out/${tag}_bench_n${N}.json; tail -3 gpurun_out/${tag}_bench_n${N}.err
for psf in gaussian7 measured31; do
  timeout 300 $TR 29521 tools/bench_bands.py --psf $psf --reps 2 > gpurun_out/${tag}_bands_${psf}_n${N}.json 2> gpurun_out/${tag}_bands_${psf}_n${N}.err; echo "bands $psf rc=$?"
  cat gpurun_out/${tag}_bands_${psf}_n${N}.json; tail -2 gpurun_out/${tag}_bands_${psf}_n${N}.err
done
SR459_BAND_NOPUSH=1 timeout 300 $TR 29522 tools/bench_bands.py --psf gaussian7 --reps 2 > gpurun_out/${tag}_bands_g7_nopush_n${N}.json 2>/dev/null; cat gpurun_out/${tag}_bands_g7_nopush_n${N}.json
timeout 300 $TR 29523 tools/bench_bands.py --psf measured7 --fractional --iters 50 --h 768 --w 1024 --reps 2 > gpurun_out/${tag}_bands_frac_n${N}.json 2> gpurun_out/${tag}_bands_frac_n${N}.err; cat gpurun_out/${tag}_bands_frac_n${N}.json; tail -2 gpurun_out/${tag}_bands_frac_n${N}.err
