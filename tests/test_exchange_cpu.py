"""World-size-2 gloo test of the multi-GPU exchange's host logic (counts exchange, variable-split all-to-all, page
reassembly, partial->final plan).  The CUDA partitioner is replaced by the oracle's CPU partition function here; the
device partition kernels are checked separately in test_partition_gpu.py."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    import oracle_lib as O
    import presto_b200 as P
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(100 + rank)
        n = 5000 + 700 * rank
        keys = rng.integers(0, 900, n).astype(np.int64)
        dates = rng.integers(9000, 9004, n).astype(np.int32)
        vals = rng.integers(-10**9, 10**9, n).astype(np.int64)
        vnull = (rng.random(n) < 0.1).astype(np.uint8) if rank == 1 else None     # only one rank has a null mask
        dec = P.Block.int128_array([int(x) * 10**20 for x in rng.integers(-50, 50, n)])
        page = P.Page(P.Block.long_array(keys), P.Block.int_array(dates, None, P.DATE), P.Block.long_array(vals, vnull, P.SHORT_DECIMAL), dec)
        ex = P.HashPartitionExchange(partitioner=O.OraclePartitioner())
        got = ex.exchange(page, [0, 1])
        # every received row belongs to this rank's partition
        ids = O.partition_ids(got.blocks[:2], world)
        assert (ids == rank).all()
        # local SINGLE-step aggregation over the received rows (oracle operators)
        gb = O.OracleGroupBy([P.BIGINT, P.DATE], 1000)
        gids = gb.get_group_ids(got.blocks[:2])
        g = gb.group_count()
        acc = O.OracleAccumulator(P.AGG_SUM, P.SHORT_DECIMAL)
        acc.add_input(gids, g, got.blocks[2])
        acc2 = O.OracleAccumulator(P.AGG_SUM, P.LONG_DECIMAL)
        acc2.add_input(gids, g, got.blocks[3])
        cnt = O.OracleAccumulator(P.AGG_COUNT)
        cnt.add_input(gids, g)
        kb, _ = gb.export(with_hash=False)
        rows = list(zip(kb[0].to_list(), kb[1].to_list(), acc.evaluate(0, g)[1][0].to_list(), acc2.evaluate(0, g)[1][0].to_list(),
                        cnt.evaluate(0, g)[1][0].to_list()))
        gathered = [None] * world
        dist.all_gather_object(gathered, (rows, page.blocks[0].values.tolist(), dates.tolist(), vals.tolist(),
                                          None if vnull is None else vnull.tolist(), dec.to_list(), got.position_count))
        if rank == 0:
            ret.put(gathered)
    finally:
        dist.destroy_process_group()


def test_exchange_world2_gloo_matches_single_process():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process expectation over the union of both ranks' inputs
    import collections
    exp_sum, exp_dec, exp_cnt = collections.defaultdict(lambda: None), collections.defaultdict(int), collections.Counter()
    total_in = 0
    for rows, keys, dates, vals, vnull, dec, n_recv in gathered:
        total_in += len(keys)
        for i, (k, d, v, x) in enumerate(zip(keys, dates, vals, dec)):
            exp_cnt[(k, d)] += 1
            exp_dec[(k, d)] += x
            if vnull is None or not vnull[i]:
                exp_sum[(k, d)] = (exp_sum[(k, d)] or 0) + v
    got = {}
    for rows, *_ in gathered:
        for k, d, s, x, c in rows:
            assert (k, d) not in got, "a group appeared on two ranks"
            got[(k, d)] = (s, x, c)
    assert sum(g[-1] for g in gathered) == total_in
    assert set(got) == set(exp_cnt)
    for key, (s, x, c) in got.items():
        assert c == exp_cnt[key] and x == exp_dec[key] and s == exp_sum[key]
