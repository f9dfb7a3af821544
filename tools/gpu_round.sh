#!/bin/bash
# one GPU-box visit: parity tests, bench (both arms), per-config timings, ncu launch lists, full ncu captures of the
# dominant kernels (k_ibp_stream for the bench line, k_lat_fwd / k_lat_bwd for the 31x31 configuration)
tag=${1:-r01x}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/${tag}_bench_ref.json
timeout 300 python tools/time_configs.py > gpurun_out/${tag}_configs.txt 2>&1; echo "configs rc=$?"
cat gpurun_out/${tag}_configs.txt
timeout 200 python tools/bench_stream.py --sets 100 --graph > gpurun_out/${tag}_stream.jsonl 2>/dev/null
timeout 200 python tools/bench_stream.py --sets 60 --fps 48 --graph >> gpurun_out/${tag}_stream.jsonl 2>/dev/null
cat gpurun_out/${tag}_stream.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/${tag}_ncu1.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ibp -s 20 -c 1 -f -o gpurun_out/${tag}_k_ibp \
    python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/${tag}_ncu2.log 2>&1; echo "ncu full rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${tag}_c4_launches.csv \
    python tools/time_configs.py c4 > gpurun_out/${tag}_ncu3.log 2>&1; echo "ncu c4 list rc=$?"
for k in k_lat_fwd k_lat_bwd; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 5 -c 1 -f -o gpurun_out/${tag}_$k \
      python tools/time_configs.py c4 > gpurun_out/${tag}_ncu4.log 2>&1; echo "ncu $k rc=$?"
done
