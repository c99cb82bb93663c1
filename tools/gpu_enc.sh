#!/bin/bash
# Encode-only check: parity tests of Encode + bench line (kernel breakdown) without the extras
python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_parity.py -m gpu -x -q -k "encode or commit_prove or host_overlapped" 2>&1 | tail -3
python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --no-commit > gpurun_out/enc_bench.json 2> gpurun_out/enc_bench.err || tail -5 gpurun_out/enc_bench.err
python -c "
import json;d=json.loads(open('gpurun_out/enc_bench.json').read().strip().splitlines()[-1]);print('s/step',d['value'],d['parity_check'][:3]);print(d['roofline_encode']['ms_per_step'], d['roofline_encode']['frac'], d['roofline_encode'].get('int',{}).get('frac'))"
