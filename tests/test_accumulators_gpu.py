"""GPU parity: GroupedAccumulator (pg_agg_*) vs the oracle for count/sum/avg/min/max over every supported type:
single step, partial->final through the intermediate layouts, masks, nulls (AggregationTestUtils.java:112-123,245-324)."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _arg_block(P, rng, t, n, null_rate=0.1):
    nulls = (rng.random(n) < null_rate).astype(np.uint8) if null_rate else None
    if t in (P.BIGINT, P.SHORT_DECIMAL):
        return P.Block.long_array(rng.integers(-10**12, 10**12, n).astype(np.int64), nulls, t)
    if t in (P.INTEGER, P.DATE):
        return P.Block.int_array(rng.integers(-10**6, 10**6, n).astype(np.int32), nulls, t)
    if t == P.DOUBLE:
        return P.Block.double_array(rng.random(n) * 1e5 + 800, nulls)
    if t == P.LONG_DECIMAL:
        return P.Block.int128_array([int(x) * 10**18 + int(y) for x, y in zip(rng.integers(-10**9, 10**9, n), rng.integers(0, 10**9, n))], nulls)
    raise AssertionError(t)


def _kinds(P):
    out = [(P.AGG_COUNT, 0)]
    for t in (P.BIGINT, P.INTEGER, P.DOUBLE, P.SHORT_DECIMAL, P.LONG_DECIMAL):
        out += [(P.AGG_COUNT_COL, t), (P.AGG_SUM, t), (P.AGG_AVG, t), (P.AGG_MIN, t), (P.AGG_MAX, t)]
    out += [(P.AGG_MIN, P.DATE), (P.AGG_MAX, P.DATE)]
    return out


def _compare(P, got, want, kind, t):
    got, want = got.to_host(), want.to_host()
    assert got.n == want.n
    gn = [got.is_null(i) for i in range(got.n)]
    wn = [want.is_null(i) for i in range(want.n)]
    assert gn == wn
    if got.type == P.DOUBLE:
        g = np.array([0.0 if x is None else x for x in got.to_list()])
        w = np.array([0.0 if x is None else x for x in want.to_list()])
        if kind in (P.AGG_MIN, P.AGG_MAX):
            assert np.array_equal(g, w, equal_nan=True)
        else:
            assert np.allclose(g, w, rtol=1e-12, atol=0), (kind, t)   # north_star: 1e-12 relative for double sums/avg
    else:
        assert got.to_list() == want.to_list(), (kind, t)


@pytest.mark.parametrize("groups,n", [(1, 5000), (4, 200000), (37, 100000), (5000, 60000)])
def test_single_step_and_partial_final(pgpu, groups, n):
    P = pgpu
    rng = np.random.default_rng(groups)
    ids = rng.integers(0, groups, n).astype(np.int32)
    ids[:groups] = np.arange(groups)
    gids = P.GroupByIdBlock(groups + 2, ids)   # two groups never touched -> null outputs
    mask = P.Block.byte_array((rng.random(n) < 0.8).astype(np.uint8))
    for kind, t in _kinds(P):
        arg = _arg_block(P, rng, t, n) if kind != P.AGG_COUNT else None
        for use_mask in (False, True):
            m = mask if use_mask else None
            acc = P.GroupedAccumulator(kind, t)
            orc = O.OracleAccumulator(kind, t)
            # two pages: host blocks then device blocks
            h = n // 2
            first = lambda b: None if b is None else P.Block(b.type, b.values[: (2 * h if b.type == P.LONG_DECIMAL else h)], None if b.nulls is None else b.nulls[:h], None, h)
            second = lambda b: None if b is None else P.Block(b.type, b.values[(2 * h if b.type == P.LONG_DECIMAL else h):], None if b.nulls is None else b.nulls[h:], None, n - h)
            acc.add_input(P.GroupByIdBlock(groups + 2, ids[:h]), first(arg), first(m))
            b2, m2 = second(arg), second(m)
            import torch
            acc.add_input(P.GroupByIdBlock(groups + 2, torch.from_numpy(ids[h:]).cuda()), None if b2 is None else b2.to_device(),
                          None if m2 is None else m2.to_device())
            assert orc.add_input(ids, groups + 2, arg, m) == 0
            rc, want, _ = orc.evaluate(0, groups + 2, True)
            assert rc == 0
            _compare(P, acc.evaluate_final(0, groups + 2), want[0], kind, t)
            # partial -> final: intermediate blocks from the GPU partial feed a GPU final and an oracle final
            inter = acc.evaluate_intermediate(0, groups + 2)
            rc, winter, _ = orc.evaluate(0, groups + 2, False)
            assert len(inter) == len(winter)
            for a, b in zip(inter, winter):
                _compare(P, a, b, kind, t)
            fin = P.GroupedAccumulator(kind, t)
            ofin = O.OracleAccumulator(kind, t)
            perm = rng.permutation(groups + 2).astype(np.int32)   # final-side group ids differ from partial-side ones
            fin.add_intermediate(P.GroupByIdBlock(groups + 2, perm), inter)
            fin.add_intermediate(P.GroupByIdBlock(groups + 2, perm), [b.to_device() for b in inter])
            ofin.add_intermediate(perm, groups + 2, winter)
            ofin.add_intermediate(perm, groups + 2, winter)
            rc, wfin, _ = ofin.evaluate(0, groups + 2, True)
            assert rc == 0
            _compare(P, fin.evaluate_final(0, groups + 2), wfin[0], kind, t)
