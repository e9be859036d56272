"""GPU parity for the f4 row of SURVEY.md §8: MarkDistinct, SetBuilder + HashSemiJoin, LookupOuter.
Golden rows are the reference's own: TestMarkDistinctOperator.java:92-110, TestHashSemiJoinOperator.java:95-431; the randomised cases
are checked against the oracle's GroupByHash (first-seen ids, contains) with the operators' rules restated in the test."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _bool_column(block):
    h = block.to_host()
    return [None if (h.nulls is not None and h.nulls[i]) else bool(h.values[i]) for i in range(h.n)]


@pytest.mark.parametrize("hash_enabled", [False, True])
def test_mark_distinct_reference_rows(pgpu, hash_enabled):
    """testMarkDistinct: two pages 0..99; every value is marked in the first page and not in the second."""
    P = pgpu
    op = P.MarkDistinctOperator([P.BIGINT], [0], hash_channel=1 if hash_enabled else None)
    got = []
    for _ in range(2):
        v = P.Block.long_array(np.arange(100, dtype=np.int64))
        blocks = [v] + ([P.Block.long_array(O.hash_rows([v]))] if hash_enabled else [])
        op.add_input(P.Page(*blocks))
        out = op.get_output()
        got += list(zip(out.blocks[0].to_host().values.tolist(), _bool_column(out.blocks[-1])))
    assert sorted(got) == sorted([(i, True) for i in range(100)] + [(i, False) for i in range(100)])


@pytest.mark.parametrize("types", ["bigint", "varchar_date"])
def test_mark_distinct_random_pages_match_oracle(pgpu, types):
    """MarkDistinctHash.markDistinctRows restated over the oracle's group ids: marked iff ids[i] == nextDistinctId (then ++), with
    repeats inside a page, across pages, and null keys."""
    P = pgpu
    rng = np.random.default_rng(31)
    kinds = [P.BIGINT] if types == "bigint" else [P.VARCHAR, P.DATE]
    op = P.MarkDistinctOperator(kinds, list(range(len(kinds))))
    ogb = O.OracleGroupBy(kinds, 100)
    next_id = 0
    for n in (5000, 1, 0, 70_000, 300):
        nulls = (rng.random(n) < 0.02).astype(np.uint8)
        if types == "bigint":
            blocks = [P.Block.long_array(rng.integers(0, 20_000, n).astype(np.int64), nulls)]
        else:
            blocks = [P.Block.variable_width(["k%d" % x for x in rng.integers(0, 300, n)]), P.Block.int_array(rng.integers(0, 40, n).astype(np.int32), nulls, P.DATE)]
        if n == 0:
            continue
        ids = ogb.get_group_ids(blocks)
        want = np.zeros(n, dtype=bool)
        for i in range(n):
            if ids[i] == next_id:
                want[i] = True
                next_id += 1
        op.add_input(P.Page(*blocks))
        got = np.array(_bool_column(op.get_output().blocks[-1]))
        assert np.array_equal(got, want)
    assert op.hash.get_group_count() == ogb.group_count()


SEMI_CASES = {
    # build rows, probe rows, expected (TestHashSemiJoinOperator.java)
    "testSemiJoin": ([10, 30, 30, 35, 36, 37, 50], list(range(30, 40)), [True, False, False, False, False, True, True, True, False, False]),
    "testBuildSideNulls": ([0, 1, 2, 2, 3, None], [1, 2, 3, 4], [True, True, True, None]),
    "testProbeSideNulls": ([0, 1, 3], [0, None, 1, 2], [True, None, True, False]),
    "testProbeAndBuildNulls": ([0, 1, None, 3], [0, None, 1, 2], [True, None, True, None]),
    "emptySetNullProbe": ([], [None, 5], [False, False]),
}


def _nullable_longs(P, rows):
    vals = np.array([0 if r is None else r for r in rows], dtype=np.int64)
    nulls = np.array([r is None for r in rows], dtype=np.uint8)
    return P.Block.long_array(vals, nulls if nulls.any() else None)


@pytest.mark.parametrize("case", sorted(SEMI_CASES))
@pytest.mark.parametrize("hash_enabled", [False, True])
def test_semi_join_reference_rows(pgpu, case, hash_enabled):
    P = pgpu
    build, probe, want = SEMI_CASES[case]
    sb = P.SetBuilderOperator(P.BIGINT, 0, hash_channel=1 if hash_enabled else None, expected_positions=10)
    if build:
        b = _nullable_longs(P, build)
        sb.add_input(P.Page(*([b] + ([P.Block.long_array(O.hash_rows([b]))] if hash_enabled else []))))
    sb.finish()
    op = P.HashSemiJoinOperator(sb, 0, probe_hash_channel=2 if hash_enabled else None)
    pb = _nullable_longs(P, probe)
    blocks = [pb, P.Block.long_array(np.arange(len(probe), dtype=np.int64))] + ([P.Block.long_array(O.hash_rows([pb]))] if hash_enabled else [])
    op.add_input(P.Page(*blocks))
    out = op.get_output()
    assert _bool_column(out.blocks[-1]) == want
    assert out.blocks[1].to_host().values.tolist() == list(range(len(probe)))      # the input channels pass through


def test_semi_join_varchar_and_large_random(pgpu):
    """testSemiJoinOnVarcharType's shape at size, rule restated over the oracle's contains()."""
    P = pgpu
    rng = np.random.default_rng(37)
    build = ["s%d" % x for x in rng.integers(0, 50_000, 120_000)]
    sb = P.SetBuilderOperator(P.VARCHAR, 0, expected_positions=10)
    for lo in range(0, len(build), 50_000):
        sb.add_input(P.Page(P.Block.variable_width(build[lo:lo + 50_000])))
    sb.finish()
    assert not sb.contains_null and sb.size() == len(set(build))
    probe = ["s%d" % x for x in rng.integers(0, 80_000, 200_000)]
    op = P.HashSemiJoinOperator(sb, 0)
    op.add_input(P.Page(P.Block.variable_width(probe)))
    got = _bool_column(op.get_output().blocks[-1])
    members = set(build)
    assert got == [p in members for p in probe]


@pytest.mark.parametrize("dups", [False, True])
def test_lookup_outer_emits_exactly_the_unvisited_build_rows(pgpu, dups):
    """OuterLookupSource + LookupOuterOperator (OuterLookupSource.java:98-160, LookupOuterOperator.java:140-181): every build position
    appended to some probe output is visited; afterwards the others come out ascending with NULL probe channels — including build
    rows with NULL keys and all rows of keys the probe never had."""
    P = pgpu
    rng = np.random.default_rng(41)
    n = 50_000
    bkeys = rng.integers(0, 30_000 if dups else 10**9, n).astype(np.int64)
    bnull = (rng.random(n) < 0.02).astype(np.uint8)
    pay = rng.integers(-10**9, 10**9, n).astype(np.int64)
    idx = P.PagesIndex([P.BIGINT, P.BIGINT], [0])
    idx.add_page([P.Block.long_array(bkeys, bnull).to_device(), P.Block.long_array(pay).to_device()])
    ls = idx.create_lookup_source()
    ls.track_visited()
    visited = np.zeros(n, dtype=bool)
    for page in range(3):
        m = 40_000
        pkeys = np.where(rng.random(m) < 0.5, bkeys[rng.integers(0, n, m)], rng.integers(0, 10**9, m)).astype(np.int64)
        pnull = (rng.random(m) < 0.02).astype(np.uint8)
        pk = P.Block.long_array(pkeys, pnull)
        if page == 1:     # the one-call form must track as well
            _, out = ls.join_and_append([pk.to_device()], [1], None, P.JOIN_INNER)
        gp, gb = ls.join_page([pk.to_device()], None, P.JOIN_PROBE_OUTER)
        gb = gb.cpu().numpy()
        visited[gb[gb >= 0]] = True
    want = np.flatnonzero(~visited)
    assert 0 < len(want) < n
    assert np.array_equal(ls.outer_positions(), want)
    op = P.LookupOuterOperator(ls, [P.BIGINT, P.DOUBLE], [0, 1], page_rows=7000)
    rows = 0
    while True:
        out = op.get_output()
        if out is None:
            break
        k = out.position_count
        pos = want[rows:rows + k]
        for c in range(2):
            assert np.asarray(out.blocks[c].to_host().nulls).all()                      # probe channels are NULL
        bk = out.blocks[2].to_host()
        assert np.array_equal(np.asarray(bk.nulls) != 0, bnull[pos] != 0)
        ok = bnull[pos] == 0
        assert np.array_equal(np.asarray(bk.values)[ok], bkeys[pos][ok])
        assert np.array_equal(np.asarray(out.blocks[3].to_host().values), pay[pos])
        rows += k
    assert rows == len(want) and op.is_finished()


def _sequence_page(P, length, start, hash_enabled):
    v = P.Block.long_array(np.arange(start, start + length, dtype=np.int64))
    return P.Page(*([v] + ([P.Block.long_array(O.hash_rows([v]))] if hash_enabled else [])))


@pytest.mark.parametrize("hash_enabled", [False, True])
@pytest.mark.parametrize("pages,limit,want", [(((3, 1), (5, 2)), 5, [1, 2, 3, 4, 5]),        # testDistinctLimit
                                              (((3, 1), (3, 2)), 3, [1, 2, 3]),              # testDistinctLimitWithPageAlignment
                                              (((3, 1), (3, 2)), 5, [1, 2, 3, 4])])          # testDistinctLimitValuesLessThanLimit
def test_distinct_limit_reference_rows(pgpu, hash_enabled, pages, limit, want):
    """TestDistinctLimitOperator.java:92-156, driven the way assertOperatorEquals drives an operator."""
    P = pgpu
    op = P.DistinctLimitOperator([P.BIGINT], [0], limit, hash_channel=1 if hash_enabled else None)
    got = []
    for length, start in pages:
        if not op.needs_input():
            break
        op.add_input(_sequence_page(P, length, start, hash_enabled))
        out = op.get_output()
        if out is not None:
            assert len(out.blocks) == (2 if hash_enabled else 1)           # the hash channel is passed through
            got += out.blocks[0].to_host().values.tolist()
    op.finish()
    assert op.is_finished() and op.get_output() is None
    assert got == want


def test_distinct_limit_random_pages_match_oracle(pgpu):
    """First row of every new group in page order, cut at the limit, with nulls and repeats; checked against the oracle's group ids."""
    P = pgpu
    rng = np.random.default_rng(5)
    limit = 7_500
    op = P.DistinctLimitOperator([P.BIGINT, P.VARCHAR], [1, 0], limit)
    ogb = O.OracleGroupBy([P.BIGINT, P.VARCHAR], 100)
    next_id, produced = 0, 0
    while op.needs_input():
        n = int(rng.integers(1, 4000))
        strings = ["s%d" % x for x in rng.integers(0, 60, n)]
        s = P.Block.variable_width(strings)
        b = P.Block.long_array(rng.integers(0, 400, n).astype(np.int64), (rng.random(n) < 0.03).astype(np.uint8))
        ids = ogb.get_group_ids([b, s])
        want = []
        for i in range(n):
            if ids[i] == next_id and produced + len(want) < limit:
                want.append(i)
                next_id += 1
        op.add_input(P.Page(s, b))
        out = op.get_output()
        produced += len(want)
        if not want:
            assert out is None
            continue
        hb = b.to_host()
        gb, gs = out.blocks[0].to_host(), out.blocks[1].to_host()
        assert np.array_equal(np.asarray(gb.nulls) != 0, np.asarray(hb.nulls)[want] != 0)
        assert np.array_equal(gb.values[np.asarray(gb.nulls) == 0], hb.values[want][np.asarray(hb.nulls)[want] == 0])
        got_strings = [bytes(gs.values[gs.offsets[i]:gs.offsets[i + 1]]).decode() for i in range(gs.n)]
        assert got_strings == [strings[i] for i in want]
        if produced > 20_000:
            break
    assert produced == limit and op.is_finished()
