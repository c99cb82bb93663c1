#!/bin/bash
# launch-shape knobs of the block pipeline (env, no rebuild): jobs per CTA of k_mac_rd and of the block transform kernels
run() { python bench.py --steps 2 --warmup 2 --no-cpu --no-e2e --no-commit --no-parity --batch 1184 > gpurun_out/knob.json 2> gpurun_out/knob.err || tail -3 gpurun_out/knob.err
  python -c "
import json;d=json.loads(open('gpurun_out/knob.json').read().strip().splitlines()[-1]);b=d['kernel_breakdown_one_step'];print('$1','s/step',round(d['value'],4),{k:round(b[k]['ms']) for k in ('k_mac','k_down','k_intt_c1','k_pintt')})"; }
run base
LMN_MAC_JB=1 run mac_jb1
LMN_MAC_JB=4 run mac_jb4
LMN_MAC_JB=8 run mac_jb8
LMN_BLK_JB=2 run blk_jb2
LMN_BLK_JB=8 run blk_jb8
LMN_BLK_JB=16 run blk_jb16
