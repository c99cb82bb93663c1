#!/bin/bash
# one config-D line on 8 GPUs (self-checked), final code
TAG=${1:-r2h}
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29708 bench.py --gpus 8 --config D --no-cpu \
  > gpurun_out/${TAG}_bench_D_8gpu.json 2> gpurun_out/${TAG}_bench_D_8gpu.err || tail -20 gpurun_out/${TAG}_bench_D_8gpu.err
python tools/benchline.py gpurun_out/${TAG}_bench_D_8gpu.json
