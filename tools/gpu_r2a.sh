#!/bin/bash
# round 2, first GPU pass on one B200: the whole -m gpu suite, smoke, the default bench line and the reference arm
set -x
nvidia-smi -L
(time python -m pytest tests -m gpu -x -q) > gpurun_out/r2a_gpu_tests.txt 2>&1; tail -6 gpurun_out/r2a_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.txt 2>&1; tail -3 gpurun_out/r2a_smoke.txt
python bench.py > gpurun_out/r2a_bench_D_1gpu.json 2> gpurun_out/r2a_bench_D_1gpu.err || tail -15 gpurun_out/r2a_bench_D_1gpu.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err || tail -5 gpurun_out/r2a_bench_ref.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2a_bench_D_1gpu.json'))
print('s/step',d['value'],'e2e',d['e2e']['value'],'parity',d['parity_check'])
print('roofline',d['roofline']['frac'],'enc',d['roofline_encode']['frac'],d['roofline_encode']['frac_two_pass'])
for k,v in d['kernel_rooflines'].items(): print(k, round(v['frac'],3), v['ms'], round(v['share_of_step'],3))
print(d['cpu_baseline']['value'], d['cpu_baseline']['parts'])
PY
