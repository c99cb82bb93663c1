#!/bin/bash
# batch-size sweep of the headline step (no extras)
for B in 296 592 888 1184 2368; do
  python bench.py --steps 2 --warmup 2 --no-cpu --no-e2e --no-commit --no-parity --batch $B > gpurun_out/batch_$B.json 2> gpurun_out/batch_$B.err || tail -3 gpurun_out/batch_$B.err
  python -c "
import json;d=json.loads(open('gpurun_out/batch_$B.json').read().strip().splitlines()[-1]);print($B,'s/step',d['value'],'roofline',round(d['roofline']['frac'],4))"
done
