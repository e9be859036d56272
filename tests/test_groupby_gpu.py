"""GPU parity: GroupByHash (pg_groupby_*) vs the oracle, re-expressing OPT/TestGroupByHash.java cases
(testAddPage :70, testNullGroup :99, testGetGroupIds :121, testAppendTo :151, testContains :205,
testContainsMultipleColumns :222, testForceRehash :238) plus seeded random pages."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _seq(n, start=0):
    return np.arange(start, start + n, dtype=np.int64)


def _both(P, types, expected, has_hash=False):
    return P.GroupByHash(types, expected, has_hash), O.OracleGroupBy(types, expected, has_hash)


def _assert_blocks_equal(a, b):
    a, b = a.to_host(), b.to_host()
    assert a.n == b.n
    if a.type == 4:   # DOUBLE: compare raw bits (NaN keys, signed zeros)
        assert [a.is_null(i) for i in range(a.n)] == [b.is_null(i) for i in range(b.n)]
        av = [0 if a.is_null(i) else int(a.values[i]) for i in range(a.n)]
        bv = [0 if b.is_null(i) else int(b.values[i]) for i in range(b.n)]
        assert av == bv
        return
    assert a.to_list() == b.to_list()


@pytest.mark.parametrize("device", [False, True])
def test_add_page_and_group_ids_first_seen_order(pgpu, device):
    P = pgpu
    gb, orc = _both(P, [P.BIGINT], 100)
    for tries in range(0, 10):   # testAddPage: re-adding the same values keeps group count; groupId == value
        blk = P.Block.long_array(_seq(10))
        blk = blk.to_device() if device else blk
        gb.add_page([blk])
        assert gb.get_group_count() == 10
        ids = gb.get_group_ids([blk])
        got = ids.ids.cpu().numpy() if device else ids.ids
        assert ids.get_group_count() == 10
        assert got.tolist() == list(range(10))
        assert orc.get_group_ids([blk]).tolist() == got.tolist()


def test_null_group(pgpu):
    P = pgpu
    gb, orc = _both(P, [P.BIGINT], 100)
    blk = P.Block.long_array([0, 0, 1], nulls=[1, 0, 0])   # testNullGroup: null then 0 must be different groups
    ids = gb.get_group_ids([blk])
    assert ids.ids.tolist() == orc.get_group_ids([blk]).tolist() == [0, 1, 2]
    assert gb.get_group_count() == 3
    keys, hashes = gb.append_values_to(with_hash=True)
    okeys, ohashes = orc.export()
    _assert_blocks_equal(keys[0], okeys[0])
    assert hashes.tolist() == ohashes.tolist()
    assert keys[0].to_list() == [None, 0, 1]


def test_append_to_and_raw_hash(pgpu):
    P = pgpu
    gb, orc = _both(P, [P.BIGINT], 100)
    vals = _seq(100)
    ids = gb.get_group_ids([P.Block.long_array(vals)])
    assert ids.ids.tolist() == list(range(100))
    orc.add_page([P.Block.long_array(vals)])
    keys, hashes = gb.append_values_to(with_hash=True)
    okeys, ohashes = orc.export()
    assert keys[0].to_list() == list(range(100))
    assert hashes.tolist() == ohashes.tolist()
    assert gb.get_raw_hash(7) == int(ohashes[7])


def test_contains(pgpu):
    P = pgpu
    gb, orc = _both(P, [P.DOUBLE], 100)
    page = [P.Block.double_array(np.arange(10, dtype=np.float64))]
    gb.add_page(page)
    orc.add_page(page)
    probe = [P.Block.double_array([3.0, 11.0, 0.0, -0.0, float("nan")])]
    assert gb.contains(probe).tolist() == orc.contains(probe).tolist() == [True, False, True, True, False]


def test_contains_multiple_columns(pgpu):
    P = pgpu
    types = [P.DOUBLE, P.VARCHAR, P.BIGINT]
    gb, orc = _both(P, types, 100)
    page = [P.Block.double_array([0.0, 1.0, 2.0]), P.Block.variable_width(["a", "bb", None]), P.Block.long_array([5, 6, 7])]
    gb.add_page(page)
    orc.add_page(page)
    probe = [P.Block.double_array([0.0, 1.0, 2.0, 2.0]), P.Block.variable_width(["a", "b", None, "x"]), P.Block.long_array([5, 6, 7, 7])]
    assert gb.contains(probe).tolist() == orc.contains(probe).tolist() == [True, False, True, False]


def test_force_rehash_small_expected_size(pgpu):
    P = pgpu
    gb, orc = _both(P, [P.BIGINT], 4)    # testForceRehash: expectedSize 4, 100 keys
    vals = _seq(100)
    gb.add_page([P.Block.long_array(vals)])
    orc.add_page([P.Block.long_array(vals)])
    assert gb.get_group_count() == 100
    assert gb.get_capacity() == orc.capacity()
    assert gb.contains([P.Block.long_array(vals)]).all()


def _random_pages(rng, types, P, n, card, null_rate):
    blocks = []
    for t in types:
        nulls = (rng.random(n) < null_rate).astype(np.uint8) if null_rate > 0 else None
        if t == P.BIGINT or t == P.SHORT_DECIMAL:
            v = rng.integers(-card, card, n).astype(np.int64)
            if t == P.BIGINT and n > 10:
                v[rng.integers(0, n)] = np.int64(-0x7fffffffffffffff)   # the table's internal empty-key sentinel pattern
            blocks.append(P.Block.long_array(v, nulls, t))
        elif t in (P.INTEGER, P.DATE):
            blocks.append(P.Block.int_array(rng.integers(-card, card, n).astype(np.int32), nulls, t))
        elif t == P.DOUBLE:
            v = rng.integers(-card, card, n).astype(np.float64) / 4
            v[rng.random(n) < 0.01] = np.nan
            v[rng.random(n) < 0.01] = -0.0
            blocks.append(P.Block.double_array(v, nulls))
        elif t == P.LONG_DECIMAL:
            ints = [int(x) * (10 ** 20) + 7 for x in rng.integers(-card, card, n)]
            blocks.append(P.Block.int128_array(ints, nulls))
        elif t == P.VARCHAR:
            strs = [None if (nulls is not None and nulls[i]) else ("k%d" % x) * (1 + x % 3) for i, x in enumerate(rng.integers(0, card, n))]
            blocks.append(P.Block.variable_width(strs))
        elif t == P.BOOLEAN:
            blocks.append(P.Block.byte_array(rng.integers(0, 2, n).astype(np.uint8), nulls))
    return blocks


CASES = [
    ("bigint", lambda P: [P.BIGINT], 3000, 0.05),
    ("bigint_dense", lambda P: [P.BIGINT], 5, 0.0),
    ("varchar1", lambda P: [P.VARCHAR], 7, 0.1),
    ("q1_keys", lambda P: [P.VARCHAR, P.VARCHAR], 3, 0.0),
    ("q3_keys", lambda P: [P.BIGINT, P.DATE, P.INTEGER], 40, 0.02),
    ("mixed", lambda P: [P.DOUBLE, P.LONG_DECIMAL, P.BOOLEAN, P.SHORT_DECIMAL], 6, 0.1),
    ("double", lambda P: [P.DOUBLE], 50, 0.05),
]


@pytest.mark.parametrize("name,types_fn,card,null_rate", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("hash_channel", [False, True])
def test_random_pages_match_oracle(pgpu, name, types_fn, card, null_rate, hash_channel):
    P = pgpu
    types = types_fn(P)
    rng = np.random.default_rng(42)
    gb, orc = _both(P, types, 16, hash_channel)
    for page_no, n in enumerate([1, 37, 4096, 20000, 0, 8192]):
        blocks = _random_pages(rng, types, P, n, card * (page_no + 1), null_rate)
        if n == 0:
            continue
        if hash_channel:
            blocks = blocks + [P.Block.long_array(O.hash_rows(blocks))]
        dev = [b.to_device() for b in blocks] if page_no % 2 else blocks
        ids = gb.get_group_ids(dev)
        got = ids.ids.cpu().numpy() if page_no % 2 else ids.ids
        want = orc.get_group_ids(blocks)
        assert np.array_equal(got, want), name
        assert ids.get_group_count() == orc.group_count()
    assert gb.get_capacity() == orc.capacity()
    keys, hashes = gb.append_values_to(with_hash=True)
    okeys, ohashes = orc.export()
    for a, b in zip(keys, okeys):
        _assert_blocks_equal(a, b)
    assert hashes.tolist() == ohashes.tolist()


def test_large_bigint_growth_and_ids(pgpu):
    """3M groups from 10M rows (BenchmarkGroupByHash.java:71-74 shape) starting from a small table: forces rebuilds."""
    P = pgpu
    rng = np.random.default_rng(7)
    vals = rng.integers(0, 3_000_000, 10_000_000).astype(np.int64)
    gb, orc = _both(P, [P.BIGINT], 1000)
    blk = P.Block.long_array(vals)
    ids = gb.get_group_ids([blk.to_device()])
    want = orc.get_group_ids([blk])
    assert np.array_equal(ids.ids.cpu().numpy(), want)
    assert ids.get_group_count() == orc.group_count()
    assert gb.get_capacity() == orc.capacity()


@pytest.mark.parametrize("kind", ["bigint", "varchar"])
def test_memory_reservation_yield(pgpu, kind):
    """TestGroupByHash.testMemoryReservationYield (TestGroupByHash.java:292-371) with UpdateMemory as a byte limit: a page far
    larger than the table's expected size; the Work blocks (process() == False, nothing consumed, and again False without more
    memory), memory is granted in steps, and at the end nothing is missing and nothing was inserted twice — for addPage and for
    getGroupIds."""
    P = pgpu
    length = 3_000_000 if kind == "bigint" else 1_200_000
    if kind == "bigint":
        values = P.Block.long_array(np.arange(length, dtype=np.int64))
        types = [P.BIGINT]
    else:
        values = P.Block.variable_width(["%d" % i for i in range(length)])
        types = [P.VARCHAR]
    page = P.Page(values)

    def drive(make_work, gb):
        limit = gb.get_estimated_size()          # what the empty table holds: any growth needs a grant
        gb.set_memory_limit(limit)
        work = make_work()
        yields = 0
        while not work.process():
            assert gb.get_group_count() == 0     # nothing consumed while blocked
            assert not work.process()            # blocked again without more memory
            assert gb.get_group_count() == 0
            yields += 1
            limit *= 2
            gb.set_memory_limit(limit)
            assert yields < 40
        return work, yields, limit

    gb = P.GroupByHash(types, 1)
    work, yields, limit = drive(lambda: gb.add_page_work(page), gb)
    assert yields >= 1
    assert gb.get_group_count() == length
    assert gb.get_estimated_size() <= limit + 64 * length      # the grant covered what the table grew to (+ per-page scratch)
    gb.close()

    gb = P.GroupByHash(types, 1)
    work, yields2, _ = drive(lambda: gb.get_group_ids_work(page), gb)
    assert yields2 == yields
    ids = work.get_result()
    assert gb.get_group_count() == length and ids.group_count == length
    got = np.asarray(ids.ids)
    assert got.shape[0] == length and np.array_equal(got, np.arange(length, dtype=np.int32))
    gb.close()
