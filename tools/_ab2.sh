# A/B of Encode variants at full config D (run under gpurun): prints the Encode kernel time per step
run() { env "$@" python bench.py --no-cpu --no-e2e --no-commit --steps 1 --warmup 1 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$*', round(d['value'],4), d['kernel_breakdown_one_step']['k_encode_pass'], d['roofline_encode']['frac'])"; }
run A=0
for f in lumenos_b200/var_*.so; do [ -f $f ] && run LUMENOS_B200_SO=/root/repo/$f; done
run LMN_ENC_CPC=8
run LMN_ENC_CPC=16
run LMN_ENC_CPC=2
