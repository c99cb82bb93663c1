#!/bin/bash
# multi-GPU parity + bench lines (run under gpurun --gpus N): TAG N [CFG:G1,G2,... ...]
TAG=$1; N=$2; shift 2
set -x
nvidia-smi -L
(time python -m pytest tests/test_gpu_multi.py -m gpu -q -rs) > gpurun_out/${TAG}_multi_tests_${N}gpu.txt 2>&1; tail -12 gpurun_out/${TAG}_multi_tests_${N}gpu.txt
for SPEC in "$@"; do          # CFG:G1,G2,...
  CFG=${SPEC%%:*}
  for G in $(echo ${SPEC#*:} | tr , ' '); do
    [ $G -le $N ] || continue
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $((29700+G)) bench.py --gpus $G --config $CFG --no-cpu \
      > gpurun_out/${TAG}_bench_${CFG}_${G}gpu.json 2> gpurun_out/${TAG}_bench_${CFG}_${G}gpu.err || tail -20 gpurun_out/${TAG}_bench_${CFG}_${G}gpu.err
    python tools/benchline.py gpurun_out/${TAG}_bench_${CFG}_${G}gpu.json
  done
done
