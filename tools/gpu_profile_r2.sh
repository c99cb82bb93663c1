#!/bin/bash
# round-2 ncu evidence on one B200: the launch list of a short config-D subset, one full capture (with source) of the dominant kernel
# and one of the other hot kernels; each only after the same command has exited 0 without ncu.  gpurun returns at most 64 MiB.
TAG=${1:-r2c}
ARGS="--cols-override 128 --batch 256 --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity"
python bench.py $ARGS > gpurun_out/${TAG}_subset_bench.json 2> gpurun_out/${TAG}_subset_bench.err || { tail -5 gpurun_out/${TAG}_subset_bench.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${TAG}_launches_raw.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_mac_rd' -s 2 -c 1 -f -o gpurun_out/${TAG}_mac python bench.py $ARGS > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none -k regex:'k_modup_blk|k_blk_fwd|k_down_pre|k_blk_intt|k_rs_chain|k_encode_pass|k_sha256_leaves|k_encrypt|k_leaf_serialize' -s 4 -c 14 -f -o gpurun_out/${TAG}_others python bench.py $ARGS > gpurun_out/${TAG}_ncu3.log 2>&1
ls -la gpurun_out/${TAG}_*
