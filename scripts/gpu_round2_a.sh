#!/bin/bash
# round 2, first GPU call: full GPU test suite, then the three bench workloads and the CPU arm
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2a_tests.log
tail -5 gpurun_out/r2a_tests.log
for wl in train seg infer; do
  python bench.py --workload $wl --steps 20 --warmup 5 --detail gpurun_out/r2a_detail_$wl.txt > gpurun_out/r2a_bench_$wl.json 2> gpurun_out/r2a_bench_$wl.err; echo "bench $wl rc=$?"
  tail -c 600 gpurun_out/r2a_bench_$wl.err
done
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2a_ref_train.json 2> gpurun_out/r2a_ref_train.err; echo "ref rc=$?"
cat gpurun_out/r2a_ref_train.json | cut -c1-400
