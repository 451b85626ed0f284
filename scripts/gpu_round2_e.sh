#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2e_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2e_tests.log
tail -8 gpurun_out/r2e_tests.log
for wl in train seg; do
  python bench.py --workload $wl --steps 20 --warmup 5 --no-cpu-baseline --detail gpurun_out/r2e_detail_$wl.txt > gpurun_out/r2e_bench_$wl.json 2> gpurun_out/r2e_bench_$wl.err; echo "bench $wl rc=$?"
  cut -c1-240 gpurun_out/r2e_bench_$wl.json
done
DC_NO_FORK=1 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2e_bench_train_nofork.json 2> gpurun_out/r2e_bench_train_nofork.err; cut -c1-240 gpurun_out/r2e_bench_train_nofork.json
