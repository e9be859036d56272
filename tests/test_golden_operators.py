"""Replay of tests/golden/reference_operator_vectors.json — inputs and expected rows transcribed from the reference's
own operator tests (OPT/TestHashJoinOperator.java, OPT/TestHashAggregationOperator.java; see tests/golden/make_golden.py).

CPU suite: the oracle must reproduce the expected rows (this is what pins the oracle at operator level).
GPU suite: the CUDA operators, through the C ABI, must reproduce the same rows (ordered for joins, like assertOperatorEquals)."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_lib as O
import presto_b200 as P

GOLDEN = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_operator_vectors.json")))
TYPES = {"varchar": P.VARCHAR, "bigint": P.BIGINT, "boolean": P.BOOLEAN}
JOIN_TYPES = {"inner": P.JOIN_INNER, "probe_outer": P.JOIN_PROBE_OUTER}
AGG_KINDS = {"count": P.AGG_COUNT, "sum": P.AGG_SUM, "avg": P.AGG_AVG, "count_col": P.AGG_COUNT_COL}


def column(type_name, values):
    if type_name == "varchar":
        return P.Block.variable_width(values)
    nulls = [v is None for v in values]
    data = np.array([0 if v is None else v for v in values], dtype=np.int64)
    if type_name == "boolean":
        return P.Block.byte_array(data.astype(np.uint8), nulls if any(nulls) else None)
    return P.Block.long_array(data, nulls if any(nulls) else None)


def page_of(type_names, rows):
    return P.Page(*[column(t, [r[c] for r in rows]) for c, t in enumerate(type_names)], position_count=len(rows))


def cell(block, i):
    v = block.get(i)
    return v.decode() if isinstance(v, bytes) else v


def materialise(probe_page, build_page, probe_idx, build_pos):
    """LookupJoinPageBuilder.build(): probe columns by index, build columns by position (-1 -> null)."""
    rows = []
    for pi, bp in zip(probe_idx, build_pos):
        row = [cell(b, int(pi)) for b in probe_page.blocks]
        row += [None if bp < 0 else cell(b, int(bp)) for b in build_page.blocks]
        rows.append(row)
    return rows


@pytest.mark.parametrize("case", GOLDEN["joins"], ids=[c["name"] for c in GOLDEN["joins"]])
def test_oracle_reproduces_reference_join_rows(case):
    build = page_of(case["build_types"], case["build"])
    probe = page_of(case["probe_types"], case["probe"])
    j = O.OracleJoin([TYPES[t] for t in case["build_types"]], [0])
    j.add_page(build)
    j.build()
    pi, bp = j.probe([probe.blocks[0]], None, JOIN_TYPES[case["join_type"]])
    assert materialise(probe, build, pi, bp) == case["expected"]


@pytest.mark.gpu
@pytest.mark.parametrize("device", [False, True])
@pytest.mark.parametrize("case", GOLDEN["joins"], ids=[c["name"] for c in GOLDEN["joins"]])
def test_gpu_reproduces_reference_join_rows(case, device):
    build = page_of(case["build_types"], case["build"])
    probe = page_of(case["probe_types"], case["probe"])
    types = [TYPES[t] for t in case["build_types"]]
    builder = P.HashBuilderOperator(types, [0])
    builder.add_input(build.to_device() if device else build)
    builder.finish()
    op = P.LookupJoinOperator(builder.get_lookup_source(), [0], list(range(len(case["probe_types"]))), list(range(len(types))),
                              JOIN_TYPES[case["join_type"]])
    op.add_input(probe.to_device() if device else probe)
    out = op.get_output()
    rows = [[cell(b.to_host(), i) for b in out.blocks] for i in range(out.position_count)]
    assert rows == case["expected"]


def agg_pages(case):
    pages = []
    for spec in case["sequence_pages"]:
        n, starts = spec[0], spec[1:]
        cols = []
        for t, s in zip(case["types"], starts):
            if t == "varchar":
                cols.append(P.Block.variable_width([str(s + i) for i in range(n)]))
            elif t == "boolean":   # BlockAssertions.createBooleanSequenceBlock: value = (i % 2 == 0)
                cols.append(P.Block.byte_array(((np.arange(s, s + n) % 2) == 0).astype(np.uint8)))
            else:
                cols.append(P.Block.long_array(np.arange(s, s + n, dtype=np.int64)))
        pages.append(P.Page(*cols))
    return pages


def check_agg_rows(case, rows):
    rows = sorted(rows)
    for r in case["expected_first_rows"]:
        assert r in rows[:50000]
    digest = hashlib.sha256(json.dumps(rows, separators=(",", ":")).encode()).hexdigest()
    assert digest == case["expected_sha256_sorted"]


@pytest.mark.parametrize("case", GOLDEN["aggregations"], ids=[c["name"] for c in GOLDEN["aggregations"]])
def test_oracle_reproduces_reference_aggregation_rows(case):
    pages = agg_pages(case)
    key_types = [TYPES[case["types"][c]] for c in case["group_by_channels"]]
    gb = O.OracleGroupBy(key_types, case["expected_groups"])
    accs = [O.OracleAccumulator(AGG_KINDS[k], 0 if ch is None else TYPES[case["types"][ch]]) for k, ch in case["aggregates"]]
    for p in pages:
        ids = gb.get_group_ids([p.blocks[c] for c in case["group_by_channels"]])
        for a, (k, ch) in zip(accs, case["aggregates"]):
            assert a.add_input(ids, gb.group_count(), None if ch is None else p.blocks[ch]) == 0
    g = gb.group_count()
    keys, _ = gb.export()
    cols = [keys[0].to_list()]
    for a in accs:
        rc, blocks, _ = a.evaluate(0, g, final=True)
        assert rc == 0
        cols.append(blocks[0].to_list())
    rows = [[c[i].decode() if isinstance(c[i], bytes) else c[i] for c in cols] for i in range(g)]
    check_agg_rows(case, rows)


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLDEN["aggregations"], ids=[c["name"] for c in GOLDEN["aggregations"]])
def test_gpu_reproduces_reference_aggregation_rows(case):
    pages = agg_pages(case)
    aggs = [P.AggregatorFactory(AGG_KINDS[k], 0 if ch is None else TYPES[case["types"][ch]], -1 if ch is None else ch)
            for k, ch in case["aggregates"]]
    op = P.HashAggregationOperator([TYPES[case["types"][c]] for c in case["group_by_channels"]], case["group_by_channels"], P.SINGLE,
                                   aggs, None, expected_groups=case["expected_groups"])
    for p in pages:
        op.add_input(p)
    op.finish()
    rows = []
    while True:
        out = op.get_output()
        if out is None:
            break
        cols = [b.to_host().to_list() for b in out.blocks]
        rows += [[c[i].decode() if isinstance(c[i], bytes) else c[i] for c in cols] for i in range(out.position_count)]
    check_agg_rows(case, rows)
