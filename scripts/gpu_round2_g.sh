#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "geometry or no_alloc or configs or hostcheck" > gpurun_out/r2g_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2g_tests.log
timeout 200 python bench_ops.py --points-log2 10 --k 10 20 30 40 --channels 64 > gpurun_out/r2g_ops_p10.jsonl 2>&1; cut -c1-200 gpurun_out/r2g_ops_p10.jsonl
timeout 200 python bench_ops.py --points-log2 20 --k 20 --channels 64 > gpurun_out/r2g_ops_big.jsonl 2>&1; cut -c1-200 gpurun_out/r2g_ops_big.jsonl
for wl in infer train; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/r2g_detail_$wl.txt > gpurun_out/r2g_bench_$wl.json 2> gpurun_out/r2g_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-200 gpurun_out/r2g_bench_$wl.json
done
