#!/bin/bash
# Round-2 multi-GPU evidence on one 8-GPU box: parity of every exchange at 2 / 4 / 8 ranks, C2 scaling lines, the other configs at
# the GPU counts BASELINE.json names, a device timeline at 8 ranks.  Usage (under gpurun --gpus 8): bash experiments/run_8gpu.sh
set -u
mkdir -p gpurun_out
O=gpurun_out
export LGB_B200_MULTI_LOG=$PWD/$O/r02_multi_gpu_parity.log
rm -f $LGB_B200_MULTI_LOG
(timeout 900 python -m pytest tests/test_gpu_multi.py -q > $O/r02_multi_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02_multi_pytest.log)
tail -3 $O/r02_multi_pytest.log
run() {  # run <N> <outfile> <bench args...>
  local n=$1 out=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
      bench.py --gpus $n "$@" > $O/$out 2> $O/${out%.json}.err
  echo "$out rc=$? $(tail -c 300 $O/$out | head -c 0)"
}
for n in 2 4 8; do run $n r02_bench_C2_${n}gpu.json --steps 20 --warmup 3; done
run 8 r02_bench_C2_8gpu_weak.json --steps 20 --warmup 3 --scaling weak --no-e2e
LGB_B200_SLICED=1 run 8 r02_bench_C2_8gpu_sliced.json --steps 20 --warmup 3 --no-e2e
LGB_B200_TRACE=$PWD/$O/r02_trace_8gpu.txt LGB_B200_TRACE_FROM=12 run 8 r02_bench_C2_8gpu_traced.json --steps 20 --warmup 3 --no-e2e --no-parity-check --no-profile
python experiments/trace_summary.py $O/r02_trace_8gpu.txt.r0 > $O/r02_timeline_8gpu.txt 2>&1
run 8 r02_bench_C3_8gpu.json --config C3 --steps 10 --warmup 3
run 8 r02_bench_C4_8gpu.json --config C4 --steps 10 --warmup 3
for n in 2 4 8; do run $n r02_bench_C5_${n}gpu.json --config C5 --steps 10 --warmup 3; done
ls -la $O | grep r02_
