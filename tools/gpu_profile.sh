#!/bin/bash
# ncu evidence for profiles/ (run under gpurun, one GPU): launch list of a short config-D subset + one full capture of the top kernels
TAG=${1:-r1d}
ARGS="--cols-override 128 --batch 256 --steps 1 --warmup 1 --no-cpu --no-e2e"
python bench.py $ARGS > gpurun_out/${TAG}_subset_bench.json 2> gpurun_out/${TAG}_subset_bench.err || { tail -5 gpurun_out/${TAG}_subset_bench.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches_raw.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_mac_rd|k_modup_blk|k_blk_fwd|k_down_pre|k_blk_intt|k_rs_pre|k_encode_pass' -c 14 -f -o gpurun_out/${TAG}_full python bench.py $ARGS > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out/${TAG}_*
