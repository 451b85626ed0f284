#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2h_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2h_tests.log
for wl in seg train; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/r2h_detail_$wl.txt > gpurun_out/r2h_bench_$wl.json 2> gpurun_out/r2h_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-200 gpurun_out/r2h_bench_$wl.json
done
grep "segmax\|(64, 3, 983040" gpurun_out/r2h_detail_seg.txt
