"""pg_partition_positions_noalloc (no device allocation: fixed static scratch) against the existing pg_partition_create path, through
the C ABI.  Collected last on purpose (file name): a failure here must not hide other tests.
Bit-exact, like every integer / index comparison in this suite.  Two calls in a row share the static scratch: the second (smaller,
other partition count) must not be disturbed by what the first left behind."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _existing(P, ctx, raw_dev, n, nparts, mode):
    """part ids, rows per partition and the inverse of dest (out[dest[i]] = i) from pg_partition_create / _ids / _scatter."""
    import torch
    N = P.native
    hb = P.Block(N.BIGINT, raw_dev, None, None, n)
    hc = hb.pg_col()
    counts = (C.c_int64 * nparts)()
    h = C.c_void_p()
    N.check(ctx.handle, N.lib().pg_partition_create(ctx.handle, None, 0, C.byref(hc), n, nparts, mode, counts, C.byref(h)))
    try:
        ids = torch.empty(max(n, 1), dtype=torch.uint8, device="cuda")
        N.check(ctx.handle, N.lib().pg_partition_ids(h, ids.data_ptr(), N.MEM_DEVICE))
        rows = P.Block(N.INTEGER, torch.arange(n, dtype=torch.int32, device="cuda"), None, None, n)
        out = P.OutBlock(N.INTEGER, n, device=True, want_nulls=False)
        ic, oc = rows.pg_col(), out.pg_col_out()
        N.check(ctx.handle, N.lib().pg_partition_scatter(h, C.byref(ic), C.byref(oc)))
        torch.cuda.synchronize()
        return ids[:n].cpu().numpy(), [int(c) for c in counts], out.values[:n].cpu().numpy()
    finally:
        N.lib().pg_partition_destroy(h)


def _noalloc(P, ctx, raw_dev, n, nparts, mode):
    import torch
    N = P.native
    ids = torch.full((max(n, 1),), 255, dtype=torch.uint8, device="cuda")
    dest = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    counts = torch.full((nparts,), -7, dtype=torch.int64, device="cuda")
    rc = N.lib().pg_partition_positions_noalloc(ctx.handle, raw_dev.data_ptr(), n, nparts, mode, ids.data_ptr(), dest.data_ptr(), counts.data_ptr())
    N.check(ctx.handle, rc)
    torch.cuda.synchronize()
    d = dest[:n].cpu().numpy()
    inv = np.empty(n, dtype=np.int32)
    inv[d] = np.arange(n, dtype=np.int32)          # out[dest[i]] = i, what pg_partition_scatter does with a row-number column
    assert len(np.unique(d)) == n                  # dest is a permutation
    return ids[:n].cpu().numpy(), counts.cpu().tolist(), inv


def test_noalloc_positions_match_pg_partition_create_on_two_calls_in_a_row(pgpu, ctx):
    import torch
    P = pgpu
    rng = np.random.default_rng(2024)
    for n, nparts, mode in ((1_000_003, 8, 0), (70_001, 5, 0), (300_000, 16, 1), (1, 3, 0)):
        raw = torch.from_numpy(rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64)).cuda()
        got = _noalloc(P, ctx, raw, n, nparts, mode)
        want = _existing(P, ctx, raw, n, nparts, mode)
        assert np.array_equal(got[0], want[0]), (n, nparts, mode)
        assert got[1] == want[1] and sum(got[1]) == n, (n, nparts, mode)
        assert np.array_equal(got[2], want[2]), (n, nparts, mode)


def test_noalloc_rejects_pages_beyond_the_fixed_scratch_and_handles_empty_pages(pgpu, ctx):
    import torch
    P = pgpu
    N = P.native
    counts = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    one = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = N.lib().pg_partition_positions_noalloc(ctx.handle, one.data_ptr(), N.NOALLOC_MAX_ROWS + 1, 4, 0, one.data_ptr(), one.data_ptr(), counts.data_ptr())
    assert rc == N.PG_E_INVALID_ARGUMENT
    assert b"at most" in N.lib().pg_last_error(ctx.handle)
    rc = N.lib().pg_partition_positions_noalloc(ctx.handle, one.data_ptr(), 10, 65, 0, one.data_ptr(), one.data_ptr(), counts.data_ptr())
    assert rc == N.PG_E_INVALID_ARGUMENT
    rc = N.lib().pg_partition_positions_noalloc(ctx.handle, None, 0, 4, 0, None, None, counts.data_ptr())
    assert rc == N.PG_OK
    torch.cuda.synchronize()
    assert counts.cpu().tolist() == [0, 0, 0, 0]
