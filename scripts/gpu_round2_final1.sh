#!/bin/bash
# round 2 final evidence on 1 GPU
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/fin_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/fin_tests.log; tail -3 gpurun_out/fin_tests.log
python __graft_entry__.py smoke > gpurun_out/fin_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/fin_smoke.log
for wl in train seg infer infer_depth2; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --detail gpurun_out/fin_detail_$wl.txt > gpurun_out/fin_bench_${wl}_1gpu.json 2> gpurun_out/fin_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-180 gpurun_out/fin_bench_${wl}_1gpu.json
done
timeout 300 python bench.py --workload train --gemm-mode 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/fin_bench_train_tf32.json 2> /dev/null
for wl in train seg infer; do
  timeout 500 python bench.py --impl reference --workload $wl --steps 20 --warmup 5 > gpurun_out/fin_ref_$wl.json 2> /dev/null; cut -c1-160 gpurun_out/fin_ref_$wl.json
done
timeout 300 python bench.py --workload fps --steps 10 --warmup 3 > gpurun_out/fin_bench_fps.json 2> /dev/null
timeout 900 python bench_ops.py --points-log2 10 12 14 16 18 20 --k 10 20 30 40 --channels 16 64 128 --out gpurun_out/fin_bench_ops.md > gpurun_out/fin_bench_ops.jsonl 2>&1; echo "ops rc=$?"
for wl in train seg infer; do
  timeout 600 ncu --cache-control none --clock-control none --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv --log-file gpurun_out/fin_ncu_step_$wl.csv python bench.py --workload $wl --steps 2 --warmup 1 --no-graph --no-cpu-baseline --sustain 0 > /dev/null 2>&1; echo "ncu $wl rc=$?"
done
