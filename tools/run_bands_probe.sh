T="python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1 --master-port"
timeout 300 $T 29521 tools/bench_bands.py --psf gaussian7 2>/dev/null | tail -1 | cut -c300-520
SR459_BAND_PUSH_KERNEL=1 timeout 300 $T 29524 tools/bench_bands.py --psf gaussian7 2>/dev/null | tail -1 | cut -c300-520
SR459_BAND_NOPUSH=1 timeout 300 $T 29522 tools/bench_bands.py --psf gaussian7 2>/dev/null | tail -1 | cut -c300-520
