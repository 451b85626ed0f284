#!/bin/bash
# round 2, second GPU call: all GPU tests (no stop at first failure), operator sweep rows, ncu counters of the new build kernels
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2b_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2b_tests.log
tail -30 gpurun_out/r2b_tests.log
python bench_ops.py --points-log2 10 --k 10 20 30 40 --channels 64 --out gpurun_out/r2b_ops_p10.md > gpurun_out/r2b_ops_p10.jsonl 2>&1
python bench_ops.py --points-log2 14 20 --k 20 --channels 64 --out gpurun_out/r2b_ops_big.md > gpurun_out/r2b_ops_big.jsonl 2>&1
cat gpurun_out/r2b_ops_p10.jsonl gpurun_out/r2b_ops_big.jsonl | cut -c1-400
python scripts/run_build_once.py 20 && ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:'fit_kernel|scale_kernel|edge_mean|cloud_mean' --csv --log-file gpurun_out/r2b_build_ncu.csv python scripts/run_build_once.py 20 > gpurun_out/r2b_build_ncu.log 2>&1
tail -12 gpurun_out/r2b_build_ncu.csv | cut -c1-300
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/r2b_detail_train.txt > gpurun_out/r2b_bench_train.json 2> gpurun_out/r2b_bench_train.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/r2b_bench_train.json
