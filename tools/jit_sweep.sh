#!/bin/bash
# A/B of the run-time specialised kernel's knobs on the Q1 bench (kernel time from the library's own event timing)
for pf in 0 1 2 4; do for vec in 4 2; do
  PRESTOGPU_JIT_PREFETCH=$pf PRESTOGPU_JIT_VEC=$vec python bench.py --steps 20 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('prefetch=$pf vec=$vec', r['kernel'], 'kernel_ms', r['launch_ms'], 'frac', r['frac'], 'step_ms', round(d['ms_per_step'],3))"
done; done
