"""The synthetic TPC-H page source (presto-1_b200/tpch.py) against the column statistics the reference ships for its TPC-H connector
(plugin/trino-tpch/src/main/resources/tpch/statistics/<sf>/*.json, extracted by tools/make_tpch_stats_golden.py into
tests/golden/tpch_stats.json): row counts, value domains (min / max) and distinct counts of every column the generator produces.
The generator is not dbgen bit for bit (io.trino.tpch:tpch is not vendored in the reference), so keys, domains and small distinct
counts are exact and the data-dependent figures are held to a stated tolerance."""
import json
import os

import numpy as np
import pytest
import torch

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tpch_stats.json")))


def _tpch():
    import importlib
    return importlib.import_module("presto_b200.tpch")


def _distinct(t):
    return int(torch.unique(t).numel())


def _close(got, want, rel):
    return abs(got - want) <= rel * want


@pytest.mark.parametrize("sf_name,sf", [("sf0.01", 0.01), ("sf1.0", 1.0)])
def test_orders_and_customer_match_the_reference_statistics(sf_name, sf):
    tpch = _tpch()
    st = GOLD[sf_name]
    od = tpch.orders(sf, device="cpu")
    oc = st["orders"]["columns"]
    assert od["orderkey"].numel() == st["orders"]["rowCount"]
    assert int(od["orderkey"].min()) == oc["o_orderkey"]["min"] and int(od["orderkey"].max()) == oc["o_orderkey"]["max"]      # dbgen's sparse keys, exactly
    assert _distinct(od["orderkey"]) == oc["o_orderkey"]["distinctValuesCount"]
    assert bool((od["orderkey"][1:] > od["orderkey"][:-1]).all())                                                            # and in key order
    assert int(od["custkey"].min()) >= oc["o_custkey"]["min"] and int(od["custkey"].max()) <= oc["o_custkey"]["max"]
    assert not bool((od["custkey"] % 3 == 0).any())                                                                           # a third of the customers has no orders
    assert _close(_distinct(od["custkey"]), oc["o_custkey"]["distinctValuesCount"], 0.01)
    assert int(od["orderdate"].min()) >= oc["o_orderdate"]["min"] and int(od["orderdate"].max()) <= oc["o_orderdate"]["max"]
    assert _close(_distinct(od["orderdate"]), oc["o_orderdate"]["distinctValuesCount"], 0.01)
    assert _distinct(od["shippriority"]) == 1 and int(od["shippriority"][0]) == oc["o_shippriority"]["min"]
    assert float(od["totalprice"].min()) >= 857.71 - 1e-9 and float(od["totalprice"].max()) <= 555285.16 + 1e-9               # the SF1 domain of o_totalprice
    assert _close(_distinct(od["totalprice"]), oc["o_totalprice"]["distinctValuesCount"], 0.03)
    cu = tpch.customer(sf, device="cpu")
    cc = st["customer"]["columns"]
    assert cu["rows"] == st["customer"]["rowCount"]
    assert int(cu["custkey"].min()) == cc["c_custkey"]["min"] and int(cu["custkey"].max()) == cc["c_custkey"]["max"]
    offs, data = cu["mktsegment_offsets"].numpy(), cu["mktsegment_bytes"].numpy()
    segments = {bytes(data[offs[i]:offs[i + 1]]).decode() for i in range(min(cu["rows"], 2000))}
    assert len(segments) == cc["c_mktsegment"]["distinctValuesCount"]
    assert min(segments) == cc["c_mktsegment"]["min"] and max(segments) == cc["c_mktsegment"]["max"]


@pytest.mark.parametrize("sf_name,sf", [("sf0.01", 0.01), ("sf1.0", 1.0)])
def test_lineitem_matches_the_reference_statistics(sf_name, sf):
    tpch = _tpch()
    st = GOLD[sf_name]["lineitem"]
    c = st["columns"]
    li = tpch.lineitem(sf, device="cpu", need=("keys", "raw", "stats"))
    assert _close(li["rows"], st["rowCount"], 0.01)                                  # 1..7 lines per order, 4 on average
    assert int(li["orderkey"].min()) == c["l_orderkey"]["min"] and int(li["orderkey"].max()) == c["l_orderkey"]["max"]
    assert _distinct(li["orderkey"]) == c["l_orderkey"]["distinctValuesCount"]
    assert bool((li["orderkey"][1:] >= li["orderkey"][:-1]).all())                   # clustered by order, as dbgen writes it
    for name, col, scale in (("l_linenumber", "linenumber", 1), ("l_quantity", "quantity", 100), ("l_discount", "discount", 100), ("l_tax", "tax", 100),
                             ("l_partkey", "partkey", 1)):
        v = li[col]
        assert int(v.min()) == round(c[name]["min"] * scale) and int(v.max()) == round(c[name]["max"] * scale), name
        assert _distinct(v) == c[name]["distinctValuesCount"], name
    ep = li["extendedprice"]                                                         # cents: quantity x the retail price formula of the part
    # the extremes dbgen reports are the ones its rows happened to hit (quantity 1 x the cheapest part, 50 x the dearest): within 0.5 %
    assert _close(int(ep.min()), round(c["l_extendedprice"]["min"] * 100), 0.005) and _close(int(ep.max()), round(c["l_extendedprice"]["max"] * 100), 0.005)
    assert _close(_distinct(ep), c["l_extendedprice"]["distinctValuesCount"], 0.03)
    sd = li["shipdate"]
    assert int(sd.min()) >= 8036 and int(sd.max()) <= 10561                          # orderdate + 1 .. orderdate + 121
    assert _close(_distinct(sd), c["l_shipdate"]["distinctValuesCount"], 0.01)
    assert {chr(x) for x in torch.unique(li["returnflag"]).tolist()} == {"A", "N", "R"}
    assert {chr(x) for x in torch.unique(li["linestatus"]).tolist()} == {"F", "O"}
    # dbgen's returnflag / linestatus rules against CURRENTDATE (1995-06-17): what makes Q1 have exactly four groups
    groups = {(chr(a), chr(b)) for a, b in torch.unique(torch.stack([li["returnflag"], li["linestatus"]], 1), dim=0).tolist()}
    assert groups == {("A", "F"), ("N", "F"), ("N", "O"), ("R", "F")}
