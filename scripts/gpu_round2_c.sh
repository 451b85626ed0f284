#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.txt
python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2c_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2c_tests.log
tail -12 gpurun_out/r2c_tests.log
python scripts/exp_build_chunk.py 2>&1 | tee gpurun_out/r2c_build_chunk.txt
python scripts/run_build_once.py 20 && ncu --cache-control none --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'fit_kernel|scale_kernel|edge_mean|cloud_mean|pack_kernel' --csv --log-file gpurun_out/r2c_build_ncu.csv python scripts/run_build_once.py 20 > gpurun_out/r2c_build_ncu.log 2>&1
echo ncu rc=$?
