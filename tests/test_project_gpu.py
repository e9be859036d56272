"""CUDA PageProcessor (csrc/project.cu) against the reference's decimal vectors and against the oracle on random pages; the fused
ScanFilterAndProject -> HashAggregation operator against oracle(process) -> oracle(GroupByHash + accumulators)."""
import numpy as np
import pytest

import oracle_lib as O
import presto_b200 as P
from presto_b200 import Expr as E
from project_cases import dec_block, dec_type, decimal_vectors, pages_equal, processors, random_page, vector_case

pytestmark = pytest.mark.gpu


def test_cuda_matches_reference_decimal_vectors(ctx):
    for case in decimal_vectors():
        page, expr = vector_case(case)
        proc = P.PageProcessor(None, [expr], ctx=ctx)
        exp = case["expected"]
        if exp == "NUMERIC_VALUE_OUT_OF_RANGE":
            with pytest.raises(P.PrestoGpuError) as ei:
                proc.process(page)
            assert ei.value.code == P.native.PG_E_NUMERIC_OVERFLOW, case
            continue
        out = proc.process(page)
        t, p, s = proc.output_types[0]
        assert (t, p, s) == (dec_type(exp["precision"]), exp["precision"], exp["scale"]), case
        assert out.blocks[0].to_list() == [exp["unscaled"]], case


@pytest.mark.parametrize("name,flt,projs", processors(), ids=[p[0] for p in processors()])
@pytest.mark.parametrize("device", [False, True], ids=["host_pages", "device_pages"])
def test_random_pages_match_oracle(ctx, name, flt, projs, device):
    rng = np.random.default_rng(abs(hash(name)) % 1000)
    gpu, orc = P.PageProcessor(flt, projs, ctx=ctx), O.OraclePageProcessor(flt, projs)
    assert gpu.output_types == orc.output_types
    for n, null_rate in ((1, 0.0), (777, 0.2), (20_000, 0.05), (4096, 0.0)):
        page = random_page(rng, n, null_rate)
        rc, want = orc.process(page)
        if rc != 0:
            with pytest.raises(P.PrestoGpuError):
                gpu.process(page.to_device() if device else page)
            continue
        got = gpu.process(page.to_device() if device else page)
        pages_equal(got, want)


def test_overflow_only_counts_for_selected_rows(ctx):
    big = 10 ** 37
    page = P.Page(dec_block([5, big * 9, 7], 38, 0), P.Block.long_array(np.array([1, 0, 1], dtype=np.int64)))
    x = E.input(0, P.LONG_DECIMAL, 38, 0)
    keep = E.input(1, P.BIGINT).eq(E.bigint(1))
    out = P.PageProcessor(keep, [x * E.decimal(9, 1, 0)], ctx=ctx).process(page)
    assert out.blocks[0].to_list() == [45, 63]
    with pytest.raises(P.PrestoGpuError) as ei:
        P.PageProcessor(None, [x * E.decimal(9, 1, 0)], ctx=ctx).process(page)
    assert ei.value.code == P.native.PG_E_NUMERIC_OVERFLOW


def _oracle_aggregate(page, key_channels, key_types, aggs):
    blocks = [b.to_host() for b in page.blocks]
    gb = O.OracleGroupBy(key_types, 1000)
    keys = [blocks[c] for c in key_channels]
    ids = gb.get_group_ids(keys) if page.position_count else np.zeros(0, dtype=np.int32)
    g = gb.group_count()
    cols = [k.to_list() for k in gb.export(with_hash=False)[0]]
    for a in aggs:
        acc = O.OracleAccumulator(a.kind, a.in_type)
        if page.position_count:
            acc.add_input(ids, g, blocks[a.input_channel] if a.input_channel >= 0 else None, blocks[a.mask_channel] if a.mask_channel >= 0 else None)
        cols.append(acc.evaluate(0, g, True)[1][0].to_list())
    return [tuple(c[i] for c in cols) for i in range(g)]


@pytest.mark.parametrize("n", [1000, 300_000])
@pytest.mark.parametrize("late_groups", [False, True])
def test_fused_filter_project_aggregate_matches_oracle(ctx, n, late_groups):
    """Q1's shape from raw columns: the filter drops rows (a group living only in rejected rows must not appear), derived long
    decimals feed sum / avg.  With late_groups a key value first shows up far behind the prefix the operator samples."""
    rng = np.random.default_rng(11 + n)
    flag = rng.integers(0, 3, n)
    if late_groups:
        flag[: n // 2] = np.minimum(flag[: n // 2], 1)
    status = rng.integers(0, 2, n)
    shipdate = rng.integers(10300, 10600, n).astype(np.int32)
    status[shipdate > 10471] = 1          # status 1 also exists below the cutoff; flag 2 / status 0 rows above the cutoff vanish
    page = P.Page(
        P.Block.variable_width([b"ANR"[f:f + 1] for f in flag]), P.Block.variable_width([b"FO"[s:s + 1] for s in status]),
        dec_block([int(v) for v in rng.integers(100, 5100, n)], 12, 2), dec_block([int(v) for v in rng.integers(90_000, 10_000_000, n)], 12, 2),
        dec_block([int(v) for v in rng.integers(0, 11, n)], 12, 2), dec_block([int(v) for v in rng.integers(0, 9, n)], 12, 2, nulls=(rng.random(n) < 0.02).astype(np.uint8)),
        P.Block.int_array(shipdate, None, P.DATE))
    qty, price = E.input(2, P.SHORT_DECIMAL, 12, 2), E.input(3, P.SHORT_DECIMAL, 12, 2)
    disc, tax = E.input(4, P.SHORT_DECIMAL, 12, 2), E.input(5, P.SHORT_DECIMAL, 12, 2)
    one = E.decimal(1, 1, 0)
    disc_price = price * (one - disc)
    charge = disc_price * (one + tax)
    flt = E.input(6, P.DATE) <= E.date(10471)
    projs = [E.input(0, P.VARCHAR), E.input(1, P.VARCHAR), qty, price, disc_price, charge, disc]
    A = P.AggregatorFactory
    aggs = [A(P.AGG_SUM, P.SHORT_DECIMAL, 2), A(P.AGG_SUM, P.SHORT_DECIMAL, 3), A(P.AGG_SUM, P.LONG_DECIMAL, 4), A(P.AGG_SUM, P.LONG_DECIMAL, 5),
            A(P.AGG_AVG, P.SHORT_DECIMAL, 2), A(P.AGG_AVG, P.LONG_DECIMAL, 5), A(P.AGG_AVG, P.SHORT_DECIMAL, 6), A(P.AGG_COUNT)]
    rc, projected = O.OraclePageProcessor(flt, projs).process(page)
    assert rc == 0
    want = _oracle_aggregate(projected, [0, 1], [P.VARCHAR, P.VARCHAR], aggs)
    for dev in (False, True):
        proc = P.PageProcessor(flt, projs, ctx=ctx)
        op = P.HashAggregationOperator([P.VARCHAR, P.VARCHAR], [0, 1], P.SINGLE, aggs, ctx=ctx, input_processor=proc)
        op.add_input(page.to_device() if dev else page)
        op.finish()
        out = op.get_output()
        got = [tuple(b.to_host().to_list()[i] for b in out.blocks) for i in range(out.position_count)]
        assert got == want          # same groups in the same first-seen order, every aggregate bit-exact
        assert op._stats()[3] == n
        op.close()
