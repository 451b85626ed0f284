#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nn.py tests/test_gpu_configs.py -q -m gpu > gpurun_out/m_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/m_tests.log
timeout 600 python bench.py --workload seg --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/m_detail_seg.txt > gpurun_out/m_bench_seg.json 2> gpurun_out/m_bench_seg.err; echo "bench seg rc=$?"; cut -c1-200 gpurun_out/m_bench_seg.json; tail -3 gpurun_out/m_bench_seg.err
grep "central_edge\|bcast\|segmax" gpurun_out/m_detail_seg.txt
