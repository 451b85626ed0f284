#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
timeout 900 python -m pytest tests -m gpu -q --maxfail=20 > gpurun_out/r2j_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2j_tests.log
timeout 300 python bench_ops.py --points-log2 10 --k 10 20 30 40 --channels 64 > gpurun_out/r2j_ops_p10.jsonl 2>&1; cut -c1-330 gpurun_out/r2j_ops_p10.jsonl
timeout 300 python bench_ops.py --points-log2 14 20 --k 20 --channels 64 > gpurun_out/r2j_ops_big.jsonl 2>&1; cut -c1-330 gpurun_out/r2j_ops_big.jsonl
for wl in train infer seg; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/r2j_detail_$wl.txt > gpurun_out/r2j_bench_$wl.json 2> gpurun_out/r2j_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-200 gpurun_out/r2j_bench_$wl.json
done
