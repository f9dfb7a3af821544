#!/bin/bash
# SAA-stage visit: parity tests, bench line, full ncu captures of the two SAA kernels
tag=${1:-r03a}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json
for k in k_resample_tile k_prefilter2d; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/${tag}_$k \
      python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/${tag}_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
