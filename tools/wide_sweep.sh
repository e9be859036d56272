#!/bin/bash
for sf in 1 2.5 5 10 20; do
  python bench.py --steps 3 --extra --extra-sf $sf --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); g=d['extra']['groupby_orderkey']; k=g['kernels_ms_per_step']
print('sf=$sf rows', g['rows'], 'groups', g['groups'], 'step_ms', g['ms_per_step'], 'wide_ms', k.get('wide_agg_jit'), 'ns/row', round(k.get('wide_agg_jit',0)*1e6/g['rows'],4), 'gather', k.get('wide_gather_kernel'), 'number', k.get('wide_number_buckets_kernel'))"
done
