#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2d_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2d_tests.log
tail -8 gpurun_out/r2d_tests.log
python bench_ops.py --points-log2 10 --k 10 20 30 40 --channels 64 --out gpurun_out/r2d_ops_p10.md > gpurun_out/r2d_ops_p10.jsonl 2>&1
python bench_ops.py --points-log2 14 20 --k 20 --channels 64 --out gpurun_out/r2d_ops_big.md > gpurun_out/r2d_ops_big.jsonl 2>&1
cat gpurun_out/r2d_ops_p10.jsonl gpurun_out/r2d_ops_big.jsonl | cut -c1-330
python scripts/run_build_once.py 20 && ncu --cache-control none --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'fit_kernel|scale_kernel|edge_mean|cloud_mean|pack_kernel' --csv --log-file gpurun_out/r2d_build_ncu.csv python scripts/run_build_once.py 20 > gpurun_out/r2d_build_ncu.log 2>&1
for wl in train seg infer; do
  python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/r2d_detail_$wl.txt > gpurun_out/r2d_bench_$wl.json 2> gpurun_out/r2d_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-240 gpurun_out/r2d_bench_$wl.json
done
python bench.py --workload train --gemm-mode 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2d_bench_train_tf32.json 2> gpurun_out/r2d_bench_train_tf32.err; cut -c1-240 gpurun_out/r2d_bench_train_tf32.json
