"""GPU parity: hash-partition kernels (pg_partition_*) vs the oracle's restatement of HashGenerator.getPartition
(MAIN/operator/HashGenerator.java:24-35) and LocalPartitionGenerator.getPartition (exchange/LocalPartitionGenerator.java:50-63)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nparts,mode", [(8, 0), (3, 0), (8, 1), (2, 1), (1, 0)])
@pytest.mark.parametrize("with_hash_channel", [False, True])
def test_partition_ids_and_stable_scatter(pgpu, ctx, nparts, mode, with_hash_channel):
    P = pgpu
    N = P.native
    rng = np.random.default_rng(nparts * 10 + mode)
    n = 50_001
    k1 = rng.integers(-10**6, 10**6, n).astype(np.int64)
    k2 = rng.integers(0, 100, n).astype(np.int32)
    k2n = (rng.random(n) < 0.05).astype(np.uint8)
    keys = [P.Block.long_array(k1), P.Block.int_array(k2, k2n)]
    payload = P.Block.double_array(rng.random(n))
    dec = P.Block.int128_array([int(x) for x in rng.integers(-10**12, 10**12, n)])
    want_ids = O.partition_ids(keys, nparts, mode)
    page = P.Page(*(keys + [payload, dec]))
    hash_channel = None
    if with_hash_channel:
        page = P.Page(*(page.blocks + [P.Block.long_array(O.hash_rows(keys))]))
        hash_channel = 4
    part, counts = P.DevicePartitioner(ctx).partition(page.to_device(), [0, 1], nparts, hash_channel, mode)
    assert counts == [int((want_ids == p).sum()) for p in range(nparts)]
    order = np.argsort(want_ids, kind="stable")     # rows keep their relative order inside a partition
    for got, src in zip(part.blocks, page.blocks):
        got = got.to_host()
        k = 2 if src.type == P.LONG_DECIMAL else 1
        exp = src.values.reshape(-1, k)[order].reshape(-1) if k == 2 else src.values[order]
        assert np.array_equal(np.asarray(got.values).view(exp.dtype), exp)
        if src.nulls is not None:
            assert np.array_equal(got.nulls, src.nulls[order])
