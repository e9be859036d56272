#!/bin/bash
for sh in "" "--extra-shuffle"; do for w in 0 1; do
  PRESTOGPU_WIDE=$w python bench.py --steps 3 --extra $sh --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); g=d['extra']['groupby_orderkey']; k=g['kernels_ms_per_step']
print('wide=$w $sh', 'step_ms', g['ms_per_step'], {a:b for a,b in list(k.items())[:6]})
for n in ('join_build','join_probe'):
    j=d['extra'][n]; print('   ',n, j['ms_per_step'], list(j['kernels_ms_per_step'].items())[:3])"
done; done
