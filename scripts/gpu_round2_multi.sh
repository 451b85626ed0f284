#!/bin/bash
# multi-GPU bench lines: N = $1 GPUs, workloads in $2.. (default train seg infer); every run under its own timeout
N=${1:-2}; shift
WLS=${@:-train seg infer}
mkdir -p gpurun_out
for wl in $WLS; do
  NCCL_DEBUG=${DC_TEST_NCCL_DEBUG:-INFO} timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --steps 20 --warmup 5 > gpurun_out/r2m_bench_${wl}_n$N.json 2> gpurun_out/r2m_bench_${wl}_n$N.err
  echo "bench $wl n=$N rc=$? stdout lines: $(wc -l < gpurun_out/r2m_bench_${wl}_n$N.json)"
  cut -c1-260 gpurun_out/r2m_bench_${wl}_n$N.json
  echo "NCCL INFO lines: $(grep -c 'NCCL INFO' gpurun_out/r2m_bench_${wl}_n$N.err)"
  grep -m3 "nranks\|Init COMPLETE" gpurun_out/r2m_bench_${wl}_n$N.err | cut -c1-200
done
