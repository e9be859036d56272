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


def test_scatter_peers_with_self_pointers(pgpu, ctx):
    """pg_partition_scatter_peers on one GPU: every 'peer' buffer is a local one, so the fused partition+store kernel can be
    checked without a second device — rows of partition p land at peer_row_offset[p] + (rank inside the partition)."""
    import torch
    P = pgpu
    N = P.native
    rng = np.random.default_rng(5)
    n, nparts = 30_011, 4
    key = P.Block.long_array(rng.integers(0, 10**6, n).astype(np.int64))
    nulls = (rng.random(n) < 0.1).astype(np.uint8)
    payload = P.Block.long_array(rng.integers(-10**9, 10**9, n).astype(np.int64), nulls)
    want_ids = O.partition_ids([key], nparts, 0)
    counts = (C.c_int64 * nparts)()
    h = C.c_void_p()
    kc = P.blocks.col_array([key])
    N.check(ctx.handle, N.lib().pg_partition_create(ctx.handle, kc, 1, None, n, nparts, 0, counts, C.byref(h)))
    try:
        offsets = [7, 0, 3, 11]                      # rows already present in each receiver's buffer
        bufs = [torch.full((offsets[p] + counts[p],), -1, dtype=torch.int64, device="cuda") for p in range(nparts)]
        nbufs = [torch.full((offsets[p] + counts[p],), 9, dtype=torch.uint8, device="cuda") for p in range(nparts)]
        pv = (C.c_void_p * nparts)(*[b.data_ptr() for b in bufs])
        pn = (C.c_void_p * nparts)(*[b.data_ptr() for b in nbufs])
        po = (C.c_int64 * nparts)(*offsets)
        col = payload.pg_col()
        N.check(ctx.handle, N.lib().pg_partition_scatter_peers(h, C.byref(col), pv, pn, po))
        torch.cuda.synchronize()
        for p in range(nparts):
            sel = want_ids == p
            assert counts[p] == int(sel.sum())
            got = bufs[p].cpu().numpy()
            assert np.all(got[:offsets[p]] == -1)                       # earlier rows untouched
            assert np.array_equal(got[offsets[p]:], payload.values[sel])   # stable order inside the partition
            assert np.array_equal(nbufs[p].cpu().numpy()[offsets[p]:], nulls[sel])
    finally:
        N.lib().pg_partition_destroy(h)


@pytest.mark.parametrize("nparts,n", [(2, 30_011), (8, 4096 * 3), (13, 50_001), (64, 9_000), (4, 1)])
def test_scatter_peers_all_columns_in_one_launch(pgpu, ctx, nparts, n):
    """pg_partition_scatter_peers_all with local 'peer' buffers: every column width (1, 4, 8, 16 bytes), with and without null
    masks, lands exactly where the per-column entry point (and the oracle's stable partition order) puts it."""
    import torch
    P = pgpu
    N = P.native
    rng = np.random.default_rng(50 + nparts)
    key = P.Block.long_array(rng.integers(0, 10**6, n).astype(np.int64))
    nulls = (rng.random(n) < 0.1).astype(np.uint8)
    cols = [key,
            P.Block.long_array(rng.integers(-10**9, 10**9, n).astype(np.int64), nulls, P.SHORT_DECIMAL),
            P.Block.int_array(rng.integers(-2**31, 2**31 - 1, n).astype(np.int32)),
            P.Block(N.BOOLEAN, rng.integers(0, 2, n).astype(np.uint8), nulls),
            P.Block.int128_array([int(x) * 10**20 * (-1 if x % 3 == 0 else 1) for x in rng.integers(1, 10**9, n)]),
            P.Block.double_array(rng.random(n))]
    want_ids = O.partition_ids([key], nparts, 0)
    counts = (C.c_int64 * nparts)()
    h = C.c_void_p()
    N.check(ctx.handle, N.lib().pg_partition_create(ctx.handle, P.blocks.col_array([key]), 1, None, n, nparts, 0, counts, C.byref(h)))
    try:
        offsets = [(3 * p) % 5 for p in range(nparts)]
        tdt = {1: torch.uint8, 4: torch.int32, 8: torch.int64, 16: torch.int64}
        widths = [8, 8, 4, 1, 16, 8]
        bufs, nbufs, pv, pn = [], [], [], []
        for c, w in zip(cols, widths):
            k = 2 if w == 16 else 1
            vb = [torch.full(((offsets[p] + counts[p]) * k,), 77, dtype=tdt[w], device="cuda") for p in range(nparts)]
            nb = [torch.full((offsets[p] + counts[p],), 9, dtype=torch.uint8, device="cuda") for p in range(nparts)] if c.nulls is not None else None
            bufs.append(vb); nbufs.append(nb)
            pv += [b.data_ptr() for b in vb]
            pn += [b.data_ptr() for b in nb] if nb else [None] * nparts
        cv = (C.c_void_p * len(pv))(*pv)
        cn = (C.c_void_p * len(pn))(*pn)
        po = (C.c_int64 * nparts)(*offsets)
        N.check(ctx.handle, N.lib().pg_partition_scatter_peers_all(h, P.blocks.col_array(cols), len(cols), cv, cn, po))
        torch.cuda.synchronize()
        for ci, (c, w) in enumerate(zip(cols, widths)):
            k = 2 if w == 16 else 1
            src = np.asarray(c.values).reshape(n, k) if k == 2 else np.asarray(c.values)
            for p in range(nparts):
                sel = want_ids == p
                assert counts[p] == int(sel.sum())
                got = bufs[ci][p].cpu().numpy()
                got = got.reshape(-1, k) if k == 2 else got
                assert np.all(got[:offsets[p]] == 77)
                assert np.array_equal(got[offsets[p]:].view(np.uint8), np.ascontiguousarray(src[sel]).view(np.uint8)), (ci, p)
                if c.nulls is not None:
                    assert np.array_equal(nbufs[ci][p].cpu().numpy()[offsets[p]:], nulls[sel])
    finally:
        N.lib().pg_partition_destroy(h)
