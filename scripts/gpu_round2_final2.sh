#!/bin/bash
# final tree of round 2, 1 GPU: whole GPU suite, smoke, the bench lines with their CPU baseline, ncu launch lists
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/fin2_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/fin2_tests.log; tail -3 gpurun_out/fin2_tests.log
python __graft_entry__.py smoke > gpurun_out/fin2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/fin2_smoke.log
for wl in train seg infer infer_depth2; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --detail gpurun_out/fin2_detail_$wl.txt > gpurun_out/fin2_bench_${wl}_1gpu.json 2> gpurun_out/fin2_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-180 gpurun_out/fin2_bench_${wl}_1gpu.json
done
timeout 300 python bench.py --workload train --gemm-mode 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/fin2_bench_train_tf32.json 2> /dev/null
for wl in train seg infer; do
  timeout 600 ncu --cache-control none --clock-control none --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv --log-file gpurun_out/fin2_ncu_step_$wl.csv python bench.py --workload $wl --steps 2 --warmup 1 --no-graph --no-cpu-baseline --sustain 0 > /dev/null 2>&1; echo "ncu $wl rc=$?"
done
