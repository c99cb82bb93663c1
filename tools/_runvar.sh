for v in 0 1 2 3; do LMN_MAC_VAR=$v python bench.py --cols-override 128 --steps 1 --warmup 1 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('var=$v', {k:v['ms'] for k,v in d['kernel_breakdown_one_step'].items() if k in ('k_mac','k_modup','k_down')})"; done
