#!/bin/bash
# A/B of the butterfly formulations inside the block kernels at full config D (run under gpurun): the default build (ALU-pinned carries)
# against -DLMN_BFLY_MADWIDE, -DLMN_BFLY_PLAINADD and both (lumenos_b200/var_*.so, built with LUMENOS_B200_SO / LUMENOS_B200_NVCC_FLAGS)
TAG=${1:-r3f}
run() { env "$@" python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu --no-commit 2>gpurun_out/${TAG}_ab.err | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);r=d['roofline'];print('$*', 's/step', round(d['value'],4), d['parity_check'][:3], {k:v['ms'] for k,v in d['kernel_breakdown_one_step'].items() if v['ms']>20}, 'ceiling', round(r['peak'],1), 'product_bfly', round(r['product_butterfly_register_resident'],1), 'fr', {k:(round(v,3) if isinstance(v,float) else v) for k,v in r['fractions'].items() if k[:3] in ('vs_','mul')})" || tail -5 gpurun_out/${TAG}_ab.err; }
run A=default
for v in mw pa both; do run LUMENOS_B200_SO=/root/repo/lumenos_b200/var_$v.so; done
