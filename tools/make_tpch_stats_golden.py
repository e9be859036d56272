"""Extract the column statistics the reference ships for TPC-H (plugin/trino-tpch/src/main/resources/tpch/statistics/<sf>/*.json:
rowCount and, per column, distinctValuesCount / min / max) for the tables and columns tpch.py generates, into
tests/golden/tpch_stats.json.  Run in the development container (the GPU box has no /root/reference):

    python tools/make_tpch_stats_golden.py
"""
import json
import os

ROOT = "/root/reference/plugin/trino-tpch/src/main/resources/tpch/statistics"
WANT = {"lineitem": ["l_orderkey", "l_partkey", "l_linenumber", "l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus",
                     "l_shipdate"],
        "orders": ["o_orderkey", "o_custkey", "o_totalprice", "o_orderdate", "o_shippriority"],
        "customer": ["c_custkey", "c_mktsegment"]}

out = {"source": "plugin/trino-tpch/src/main/resources/tpch/statistics/{sf0.01,sf1.0}/{lineitem,orders,customer}.json"}
for sf in ("sf0.01", "sf1.0"):
    out[sf] = {}
    for table, cols in WANT.items():
        with open(os.path.join(ROOT, sf, table + ".json")) as f:
            st = json.load(f)
        out[sf][table] = {"rowCount": st["rowCount"], "columns": {c: {k: st["columns"][c][k] for k in ("distinctValuesCount", "min", "max")} for c in cols}}
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "tpch_stats.json")
with open(dst, "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
print("wrote", dst)
