# usage: bash tools/_runvar.sh "<env assignments>" ... ; runs a 128-column config-D bench per variant and prints the kernel breakdown
for v in "$@"; do env $v python bench.py --cols-override 128 --steps 1 --warmup 1 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$v', round(d['value']*1e3,1), {k:v['ms'] for k,v in d['kernel_breakdown_one_step'].items() if v['ms']>1})"; done
