#!/bin/bash
# after the pair-GEMM barrier fix: GEMM tests + soak of the captured inference / training steps
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nn.py -q -m gpu > gpurun_out/l_tests_nn.log 2>&1; echo "nn tests rc=$?"; tail -4 gpurun_out/l_tests_nn.log
timeout 300 python scripts/stress_gemm.py > gpurun_out/l_stress_gemm.log 2>&1; echo "stress rc=$?"; tail -3 gpurun_out/l_stress_gemm.log
timeout 200 python scripts/diag_soak.py --workload infer --mode graph --seconds 90 --rank 0 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/l_soak_infer_r0.txt
timeout 200 python scripts/diag_soak.py --workload infer --mode graph --seconds 45 --rank 1 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/l_soak_infer_r1.txt
timeout 200 python scripts/diag_soak.py --workload train --mode graph --seconds 45 --rank 0 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/l_soak_train.txt
timeout 200 python scripts/diag_soak.py --workload seg --mode graph --seconds 30 --rank 0 2>&1 | grep -a "SOAK\|Error" | tee gpurun_out/l_soak_seg.txt
timeout 300 python bench.py --workload infer --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/l_bench_infer.json 2> gpurun_out/l_bench_infer.err; echo "bench infer rc=$?"; cut -c1-200 gpurun_out/l_bench_infer.json
timeout 300 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/l_bench_train.json 2> gpurun_out/l_bench_train.err; echo "bench train rc=$?"; cut -c1-200 gpurun_out/l_bench_train.json
