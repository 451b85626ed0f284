#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_api_corners.py -m gpu -q -x > gpurun_out/r2i_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2i_tests.log
timeout 400 python bench_ops.py --morton --points-log2 10 14 20 --k 20 --channels 64 128 --out gpurun_out/r2i_ops_morton.md > gpurun_out/r2i_ops_morton.jsonl 2>&1; cut -c1-330 gpurun_out/r2i_ops_morton.jsonl
for wl in train infer seg; do
  timeout 300 python bench.py --workload $wl --morton --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2i_bench_${wl}_morton.json 2> gpurun_out/r2i_bench_${wl}_morton.err; echo "bench $wl morton rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r2i_bench_${wl}_morton.json')); print(d['value'], d['ms_per_step'], {k:v['ms_per_step'] for k,v in d['kernels'].items()})"
done
