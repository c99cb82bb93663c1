#!/bin/bash
# round 2, GPU pass d on one B200: whole -m gpu suite, smoke, bench line + reference arm, and a full ncu capture of the kernels the
# first capture missed (rescale chain, Encode passes, SHA-256 leaves, encryptor / decryptor, leaf serialiser)
TAG=r2d
set -x
(time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_gpu_tests.txt 2>&1; tail -6 gpurun_out/${TAG}_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.txt 2>&1; tail -3 gpurun_out/${TAG}_smoke.txt
python bench.py > gpurun_out/${TAG}_bench_D_1gpu.json 2> gpurun_out/${TAG}_bench_D_1gpu.err || tail -15 gpurun_out/${TAG}_bench_D_1gpu.err
python tools/benchline.py gpurun_out/${TAG}_bench_D_1gpu.json
ARGS="--cols-override 128 --batch 256 --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity"
ncu --set full --clock-control none -k regex:'k_rs_chain|k_encode_pass|k_sha256_leaves|k_encrypt|k_decrypt_limb|k_crt_decode|k_leaf_serialize|k_mul_pt|k_ct_stream' -c 12 -f -o gpurun_out/${TAG}_others python bench.py $ARGS > gpurun_out/${TAG}_ncu.log 2>&1
ls -la gpurun_out/${TAG}_*
