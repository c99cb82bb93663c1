#!/bin/bash
# round-end GPU pass on one B200: whole -m gpu suite, smoke, the default bench line, the reference arm, then the ncu launch list of the
# profiling subset and one full capture (with source) of the dominant kernel with the final code
TAG=${1:-r2g}
set -x
(time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_gpu_tests.txt 2>&1; tail -6 gpurun_out/${TAG}_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.txt 2>&1; tail -3 gpurun_out/${TAG}_smoke.txt
python bench.py > gpurun_out/${TAG}_bench_D_1gpu.json 2> gpurun_out/${TAG}_bench_D_1gpu.err || tail -15 gpurun_out/${TAG}_bench_D_1gpu.err
python tools/benchline.py gpurun_out/${TAG}_bench_D_1gpu.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err || tail -5 gpurun_out/${TAG}_bench_ref.err
ARGS="--cols-override 128 --batch 256 --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --no-commit"
python bench.py $ARGS > gpurun_out/${TAG}_subset_bench.json 2> gpurun_out/${TAG}_subset_bench.err || { tail -5 gpurun_out/${TAG}_subset_bench.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${TAG}_launches_raw.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_mac_rd' -s 2 -c 1 -f -o gpurun_out/${TAG}_mac python bench.py $ARGS > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none -k regex:'k_blk_fwd|k_blk_intt|k_rs_chain' -s 3 -c 6 -f -o gpurun_out/${TAG}_blk python bench.py $ARGS > gpurun_out/${TAG}_ncu3.log 2>&1
ls -la gpurun_out/${TAG}_*
