#!/bin/bash
# final multi-GPU evidence on one 8xB200 node: the whole multi-GPU parity file, then bench lines (each self-checked against the oracle on every
# rank): config D at 2/4/8, config C at 2/4/8 (BASELINE configs[2]), config A at 1/8 (the latency regime)
TAG=${1:-r2f}
set -x
nvidia-smi -L | wc -l
(time python -m pytest tests/test_gpu_multi.py -m gpu -q -rs) > gpurun_out/${TAG}_multi_tests_8gpu.txt 2>&1; tail -6 gpurun_out/${TAG}_multi_tests_8gpu.txt
run() { CFG=$1; G=$2
  if [ $G -eq 1 ]; then python bench.py --config $CFG --no-cpu > gpurun_out/${TAG}_bench_${CFG}_${G}gpu.json 2> gpurun_out/${TAG}_bench_${CFG}_${G}gpu.err || tail -20 gpurun_out/${TAG}_bench_${CFG}_${G}gpu.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $((29700+G)) bench.py --gpus $G --config $CFG --no-cpu \
      > gpurun_out/${TAG}_bench_${CFG}_${G}gpu.json 2> gpurun_out/${TAG}_bench_${CFG}_${G}gpu.err || tail -20 gpurun_out/${TAG}_bench_${CFG}_${G}gpu.err; fi
  python tools/benchline.py gpurun_out/${TAG}_bench_${CFG}_${G}gpu.json; }
run D 8; run D 4; run D 2; run C 8; run C 4; run C 2; run A 8; run A 1
