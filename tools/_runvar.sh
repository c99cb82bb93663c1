python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -5
for v in ra r16; do LMN_MAC_CORE=$v python bench.py --cols-override 128 --steps 1 --warmup 1 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('core=$v', {k:v['ms'] for k,v in d['kernel_breakdown_one_step'].items() if k in ('k_mac','k_modup','k_down')})"; done
