# A/B of library variants / env knobs on a 128-column config-D subset (run under gpurun)
run() { env "$@" python bench.py --cols-override 128 --steps 1 --warmup 1 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$*', round(d['value']*1e3,1), {k:v['ms'] for k,v in d['kernel_breakdown_one_step'].items() if v['ms']>1}, 'cm', {k:v['ms'] for k,v in d['commit_merkle']['kernels'].items() if v['ms']>1}, 'peak', d['roofline_int']['peak'])"; }
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -3
run A=0
for f in lumenos_b200/var_*.so; do [ -f $f ] && run LUMENOS_B200_SO=/root/repo/$f; done
