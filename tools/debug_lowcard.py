"""Debug helper: the late / unpackable varchar key scenario of tests/test_operators_gpu.py, printing rows that differ from the oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np
import presto_b200 as P
import oracle_lib as O

rng = np.random.default_rng(9)
n = 1_300_000
late = n - 200_000
base = np.array(["A", "N", "R", "", "xy"], dtype=object)
ks = base[rng.integers(0, 5, n)]
ks[late:late + 50_000] = "late"
ks[late + 60_000:late + 70_000] = "Q"
ks[rng.random(n) < 0.001] = None
key_blocks = [P.Block.variable_width(list(ks)), P.Block.variable_width(list(base[rng.integers(0, 2, n)]))]
types = [P.VARCHAR, P.VARCHAR]
nk = 2
v_short = rng.integers(-10**10, 10**10, n).astype(np.int64)
v_nulls = (rng.random(n) < 0.05).astype(np.uint8)
mask = (rng.random(n) < 0.9).astype(np.uint8)
blocks = key_blocks + [P.Block.long_array(v_short, v_nulls, P.SHORT_DECIMAL), P.Block.byte_array(mask)]
c = nk
aggs = [P.AggregatorFactory(P.AGG_COUNT, 0, -1, c + 1), P.AggregatorFactory(P.AGG_SUM, P.SHORT_DECIMAL, c, c + 1),
        P.AggregatorFactory(P.AGG_COUNT_COL, P.SHORT_DECIMAL, c), P.AggregatorFactory(P.AGG_COUNT)]
op = P.HashAggregationOperator(types, [0, 1], P.SINGLE, aggs)
op.add_input(P.Page(*blocks).to_device())
op.finish()
out = op.get_output(max_rows=1000)
got = [tuple(b.to_host().to_list()[i] for b in out.blocks) for i in range(out.position_count)]
ogb = O.OracleGroupBy(types, 10_000)
ids = ogb.get_group_ids(blocks[:nk])
g = ogb.group_count()
keys, _ = ogb.export(with_hash=False)
cols = [k.to_list() for k in keys]
for a in aggs:
    oa = O.OracleAccumulator(a.kind, a.in_type)
    oa.add_input(ids, g, blocks[a.input_channel] if a.input_channel >= 0 else None, blocks[a.mask_channel] if a.mask_channel >= 0 else None)
    cols.append(oa.evaluate(0, g, True)[1][0].to_list())
want = [tuple(col[i] for col in cols) for i in range(g)]
print("groups", len(got), len(want))
for i, (a, b) in enumerate(zip(got, want)):
    print(i, "OK " if a == b else "BAD", a, b)
