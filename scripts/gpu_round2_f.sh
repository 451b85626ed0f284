#!/bin/bash
# round 2 evidence run (1 GPU): full GPU test suite, the three bench workloads with their CPU legs, the reference arm, the
# operator sweep, ncu launch lists with DRAM traffic of one eager step per workload, ncu --set full of the fit kernel
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2f_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2f_tests.log
tail -4 gpurun_out/r2f_tests.log
for wl in train seg infer; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --detail gpurun_out/r2f_detail_$wl.txt > gpurun_out/r2f_bench_$wl.json 2> gpurun_out/r2f_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-200 gpurun_out/r2f_bench_$wl.json
done
timeout 300 python bench.py --workload train --gemm-mode 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2f_bench_train_tf32.json 2> /dev/null
timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2f_ref_train.json 2> gpurun_out/r2f_ref_train.err; echo "ref train rc=$?"; cut -c1-200 gpurun_out/r2f_ref_train.json
timeout 300 python bench.py --impl reference --workload seg --steps 20 --warmup 5 > gpurun_out/r2f_ref_seg.json 2> /dev/null; cut -c1-200 gpurun_out/r2f_ref_seg.json
timeout 300 python bench.py --impl reference --workload infer --steps 20 --warmup 5 > gpurun_out/r2f_ref_infer.json 2> /dev/null; cut -c1-200 gpurun_out/r2f_ref_infer.json
timeout 300 python bench.py --workload fps --steps 10 --warmup 3 > gpurun_out/r2f_bench_fps.json 2> /dev/null
timeout 900 python bench_ops.py --points-log2 10 12 14 16 18 20 --k 10 20 30 40 --channels 16 64 128 --out gpurun_out/r2f_bench_ops.md > gpurun_out/r2f_bench_ops.jsonl 2>&1; echo "ops rc=$?"
for wl in train seg infer; do
  timeout 600 ncu --cache-control none --clock-control none --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv --log-file gpurun_out/r2f_ncu_step_$wl.csv python bench.py --workload $wl --steps 2 --warmup 1 --no-graph --no-cpu-baseline --sustain 0 > /dev/null 2>&1; echo "ncu $wl rc=$?"
done
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'fit_kernel' -c 1 -o gpurun_out/r2f_fit_full python scripts/run_build_once.py 20 > /dev/null 2>&1; echo "ncu full fit rc=$?"
ls -la gpurun_out | tail -5
