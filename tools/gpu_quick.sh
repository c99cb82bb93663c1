#!/bin/bash
# quick GPU check after a kernel change: parity tests + a short bench D with the kernel breakdown (run under gpurun)
python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 2 --warmup 2 --no-cpu --no-e2e ${BENCH_ARGS} > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err || tail -5 gpurun_out/quick_bench.err
python -c "
import json;d=json.loads(open('gpurun_out/quick_bench.json').read().strip().splitlines()[-1]);print('s/step',d['value'],d['parity_check'][:3], 'merkle', (d.get('commit_merkle') or {}).get('ms'));print({k:v['ms'] for k,v in d['kernel_breakdown_one_step'].items()})"
