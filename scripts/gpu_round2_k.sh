#!/bin/bash
# deep centralized edge MLP (seg) + soak test of the inference replay
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nn.py -q -m gpu -k "segmax or deep_centralized or centralized" > gpurun_out/k_tests_nn.log 2>&1; echo "nn tests rc=$?"; tail -15 gpurun_out/k_tests_nn.log
timeout 600 python -m pytest tests/test_gpu_configs.py -q -m gpu > gpurun_out/k_tests_cfg.log 2>&1; echo "config tests rc=$?"; tail -5 gpurun_out/k_tests_cfg.log
timeout 600 python bench.py --workload seg --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/k_detail_seg.txt > gpurun_out/k_bench_seg.json 2> gpurun_out/k_bench_seg.err; echo "bench seg rc=$?"; cut -c1-200 gpurun_out/k_bench_seg.json; tail -3 gpurun_out/k_bench_seg.err
timeout 200 python scripts/diag_soak.py --workload infer --mode graph --seconds 60 --rank 0 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/k_soak_infer_r0.txt
timeout 200 python scripts/diag_soak.py --workload infer --mode graph --seconds 40 --rank 1 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/k_soak_infer_r1.txt
if grep -q '"status": "clean"' gpurun_out/k_soak_infer_r0.txt && grep -q '"status": "clean"' gpurun_out/k_soak_infer_r1.txt; then
  echo "soak clean"
else
  timeout 300 python scripts/diag_soak.py --workload infer --mode eager-sync --seconds 120 --rank 0 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/k_soak_infer_sync.txt
fi
nvidia-smi -q | grep -i -A3 "xid\|ecc errors" | head -30 > gpurun_out/k_smi.txt 2>&1
