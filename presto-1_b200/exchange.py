"""Hash-partitioned exchange between the GPUs of one node — the device replacement for the reference's
PartitionedOutputOperator -> OutputBuffer -> HTTP -> ExchangeOperator shuffle (MAIN/operator/PartitionedOutputOperator.java,
partition function MAIN/operator/HashGenerator.java:24-35) and of LocalExchange (MAIN/operator/exchange/).

One process per GPU; torch.distributed (NCCL over NVLink / NVSwitch) is the plumbing:
  1. pg_partition_create  : CUDA kernels compute each row's partition and its slot in partition-major order
  2. pg_partition_scatter : every column is rewritten partition-major on the device (raw column buffers, no serde)
  3. all_to_all_single    : counts first (world x world int64), then one variable-split all-to-all per column buffer
  3'. peer mode (NVLink / NVSwitch peer memory, the default under NCCL when symmetric memory is available): the receive buffers are
     symmetric-memory allocations; after the counts are known to everyone, pg_partition_scatter_peers writes every row of a column
     straight into its destination rank's buffer (one kernel does the partitioning store AND the transfer, no staging copy and
     no collective on the data path); two stream-ordered barriers bracket the stores.  The result layout is the one
     all_to_all_single gives: source-rank major, page order inside a source.
Rows with equal keys meet on one rank, so a SINGLE-step aggregation / a local build+probe per rank gives disjoint results.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .blocks import Block, OutBlock, Page, col_array
from .context import default_context

_TORCH_DT = None


def _torch_dtype(t):
    import torch
    return {N.BIGINT: torch.int64, N.DOUBLE: torch.int64, N.SHORT_DECIMAL: torch.int64, N.LONG_DECIMAL: torch.int64,
            N.INTEGER: torch.int32, N.DATE: torch.int32, N.BOOLEAN: torch.uint8}[t]


class DevicePartitioner:
    """The product path: CUDA partition + scatter (fails loudly without the library / a GPU)."""

    def __init__(self, ctx=None):
        self.ctx = ctx or default_context()

    def partition(self, page, key_channels, nparts, hash_channel=None, mode=0):
        """Returns (partition-major device page, counts per partition)."""
        import torch
        blocks = page.blocks
        n = page.position_count
        keys = [blocks[c] for c in key_channels]
        hc = blocks[hash_channel].pg_col() if hash_channel is not None else None
        counts = (C.c_int64 * nparts)()
        h = C.c_void_p()
        N.check(self.ctx.handle, N.lib().pg_partition_create(self.ctx.handle, col_array(keys), len(keys), C.byref(hc) if hc else None, n,
                                                             nparts, mode, counts, C.byref(h)))
        try:
            out_blocks = []
            for b in blocks:
                if b.type in (N.VARCHAR, N.VARBINARY):
                    raise N.PrestoGpuError(N.PG_E_INVALID_ARGUMENT, "variable width columns cannot cross the device exchange")
                o = OutBlock(b.type, n, device=True, want_nulls=b.nulls is not None)
                oc = o.pg_col_out()
                ic = b.pg_col()
                N.check(self.ctx.handle, N.lib().pg_partition_scatter(h, C.byref(ic), C.byref(oc)))
                out_blocks.append(o.block(n))
        finally:
            N.lib().pg_partition_destroy(h)
        return Page(*out_blocks, position_count=n), [int(c) for c in counts]


class _PeerBuffers:
    """Symmetric-memory receive buffers, one per (column slot, kind); grown geometrically, reused across exchanges."""

    def __init__(self, group, device):
        import torch.distributed._symmetric_memory as symm
        self.symm, self.group, self.device = symm, group, device
        self.bufs = {}

    def get(self, key, nbytes):
        import torch
        cur = self.bufs.get(key)
        if cur is None or cur[0].numel() < nbytes:
            cap = max(int(nbytes * 1.25) + 4096, 1 << 20)
            cap = (cap + 255) // 256 * 256
            t = self.symm.empty(cap, dtype=torch.uint8, device=self.device)
            hdl = self.symm.rendezvous(t, self.group)
            cur = (t, hdl)
            self.bufs[key] = cur
        return cur


class HashPartitionExchange:
    """exchange(page) -> the rows of every rank whose key hashes to this rank."""

    MAX_COLS_PER_LAUNCH = 16

    def __init__(self, partitioner=None, group=None, device=None, peer="auto"):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.partitioner = partitioner or DevicePartitioner()
        self.backend = dist.get_backend(group)
        self.device = device if device is not None else ("cuda" if self.backend == "nccl" else "cpu")
        self.sent_rows = 0
        self.received_rows = 0
        self.bytes_sent = 0
        # peer mode: decided once, by everyone together (a rank that cannot set up symmetric memory takes all ranks to NCCL)
        self.peer = None
        self.peer_why = "not requested"
        import os as _os
        self.per_column = _os.environ.get("PRESTOGPU_EXCHANGE_PER_COLUMN", "") == "1"     # the one-kernel-per-column form (kept for comparison)
        if peer and self.backend == "nccl" and self.world > 1:
            import os
            import torch
            ok = 0
            if os.environ.get("PRESTOGPU_EXCHANGE", "") != "nccl":
                try:
                    pg = group if group is not None else dist.group.WORLD
                    pb = _PeerBuffers(pg, torch.device("cuda", torch.cuda.current_device()))
                    pb.get(("probe", 0), 1024)[1].barrier()
                    torch.cuda.synchronize()
                    ok = 1
                except Exception as e:     # noqa: BLE001 - any failure means "use the collective"
                    self.peer_why = "symmetric memory unavailable: %s" % (str(e).splitlines()[0][:200] if str(e) else type(e).__name__)
            else:
                self.peer_why = "PRESTOGPU_EXCHANGE=nccl"
            t = torch.tensor([ok], dtype=torch.int64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
            if int(t.item()) == 1:
                self.peer = pb
                self.peer_why = "symmetric memory"
            elif ok:
                self.peer_why = "another rank could not set up symmetric memory"

    def _exchange_peer(self, page, key_channels, hash_channel, slot, lease):
        """Fused partition + transfer: rows are stored directly into the destination ranks' symmetric receive buffers.
        The receive buffers belong to `slot`: a page returned with lease=True is a view into them and stays valid until the next
        exchange with the same slot (on any rank); with lease=False the rows are copied out before returning."""
        import torch
        dist, W, me = self.dist, self.world, self.rank
        blocks = page.blocks
        n = page.position_count
        ctx = self.partitioner.ctx
        # the two barriers below are issued on torch's current stream and the stores on the context's stream: they must be the same
        cur = torch.cuda.current_stream().cuda_stream
        if getattr(ctx, "stream", cur) != cur:
            raise N.PrestoGpuError(N.PG_E_ILLEGAL_STATE, "peer exchange: the partitioner's context is bound to another stream than torch's current one")
        keys = [blocks[c] for c in key_channels]
        hc = blocks[hash_channel].pg_col() if hash_channel is not None else None
        counts = (C.c_int64 * W)()
        h = C.c_void_p()
        N.check(ctx.handle, N.lib().pg_partition_create(ctx.handle, col_array(keys), len(keys), C.byref(hc) if hc else None, n, W, 0, counts,
                                                        C.byref(h)))
        try:
            # everyone learns the whole counts matrix and which columns carry nulls anywhere: one small all-gather
            mine = torch.tensor([int(c) for c in counts] + [1 if b.nulls is not None else 0 for b in blocks], dtype=torch.int64, device=self.device)
            allv = torch.empty(W * mine.numel(), dtype=torch.int64, device=self.device)
            dist.all_gather_into_tensor(allv, mine, group=self.group)
            m = allv.view(W, -1).cpu().numpy()
            cm = m[:, :W]                                        # cm[s][r] rows rank s sends to rank r
            nulls_any = m[:, W:].max(axis=0) > 0
            recv_total = cm.sum(axis=0)                          # rows each rank receives
            cap_rows = int(recv_total.max())
            n_out = int(recv_total[me])
            offs = (C.c_int64 * W)(*[int(cm[:me, r].sum()) for r in range(W)])   # where my rows start in rank r's buffer
            out_blocks = []
            pending_views = []
            srcs, pv, pn, handle = [], [], [], None
            for ci, (b, nullable) in enumerate(zip(blocks, nulls_any)):
                if b.type in (N.VARCHAR, N.VARBINARY):
                    raise N.PrestoGpuError(N.PG_E_INVALID_ARGUMENT, "variable width columns cannot cross the device exchange")
                dt = _torch_dtype(b.type)
                k = 2 if b.type == N.LONG_DECIMAL else 1
                esz = torch.empty(0, dtype=dt).element_size() * k
                vbuf, vh = self.peer.get((slot, ci, "v"), max(cap_rows, 1) * esz)
                nbuf, nh = (self.peer.get((slot, ci, "n"), max(cap_rows, 1)) if nullable else (None, None))
                handle = vh
                pv += [int(vh.buffer_ptrs[r]) for r in range(W)]
                pn += [int(nh.buffer_ptrs[r]) if nullable else None for r in range(W)]
                src = b
                if nullable and b.nulls is None:     # someone else's rows have nulls: mine travel with an all-zero mask
                    src = Block(b.type, b.values, torch.zeros(n, dtype=torch.uint8, device=self.device), None, n, b.precision, b.scale)
                srcs.append(src)
                vals = vbuf[: n_out * esz].view(dt)
                nulls = nbuf[:n_out] if nullable else None
                pending_views.append(len(out_blocks))
                out_blocks.append(Block(b.type, vals, nulls, None, n_out, b.precision, b.scale))
                self.bytes_sent += (n - int(counts[me])) * esz
            if handle is not None:
                handle.barrier()          # stream-ordered: every rank's earlier readers of its receive buffers are done
                if self.per_column:
                    for ci, src in enumerate(srcs):
                        col = src.pg_col()
                        cv = (C.c_void_p * W)(*pv[ci * W:(ci + 1) * W])
                        cn = (C.c_void_p * W)(*pn[ci * W:(ci + 1) * W]) if pn[ci * W] is not None else None
                        N.check(ctx.handle, N.lib().pg_partition_scatter_peers(h, C.byref(col), cv, cn, offs))
                else:
                    for c0 in range(0, len(srcs), self.MAX_COLS_PER_LAUNCH):      # one launch stores every column of the page
                        part = srcs[c0:c0 + self.MAX_COLS_PER_LAUNCH]
                        cv = (C.c_void_p * (W * len(part)))(*pv[c0 * W:(c0 + len(part)) * W])
                        cn = (C.c_void_p * (W * len(part)))(*pn[c0 * W:(c0 + len(part)) * W])
                        N.check(ctx.handle, N.lib().pg_partition_scatter_peers_all(h, col_array(part), len(part), cv, cn, offs))
                handle.barrier()          # every rank's stores have landed before anyone reads its buffer
            if not lease:                 # the caller may hold this page across later exchanges: give it its own memory
                for i in pending_views:
                    ob = out_blocks[i]
                    out_blocks[i] = Block(ob.type, ob.values.clone(), None if ob.nulls is None else ob.nulls.clone(), None, ob.n, ob.precision, ob.scale)
        finally:
            N.lib().pg_partition_destroy(h)
        self.sent_rows += n - int(counts[me])
        self.received_rows += n_out
        return Page(*out_blocks, position_count=n_out)

    def _as_tensor(self, arr):
        import torch
        if type(arr).__module__.startswith("torch"):
            return arr.to(self.device)
        a = np.ascontiguousarray(arr)
        if a.dtype == np.uint64:
            a = a.view(np.int64)
        return torch.from_numpy(a).to(self.device)

    def exchange(self, page, key_channels, hash_channel=None, slot="default", lease=False):
        """The rows of every rank whose key hashes to this rank (source-rank major, page order inside a source).
        slot / lease only matter in peer mode: see _exchange_peer.  The NCCL path always returns fresh tensors."""
        import torch
        if self.peer is not None:
            return self._exchange_peer(page, key_channels, hash_channel, slot, lease)
        dist, W = self.dist, self.world
        part_page, counts = self.partitioner.partition(page, key_channels, W, hash_channel)
        send = torch.tensor(counts, dtype=torch.int64, device=self.device)
        recv = torch.empty(W, dtype=torch.int64, device=self.device)
        dist.all_to_all_single(recv, send, group=self.group)
        # which columns carry null masks anywhere in the job (one tiny all-reduce so that every rank agrees)
        has_nulls = torch.tensor([1 if b.nulls is not None else 0 for b in part_page.blocks], dtype=torch.int64, device=self.device)
        dist.all_reduce(has_nulls, op=dist.ReduceOp.MAX, group=self.group)
        recv_counts = [int(x) for x in recv.cpu().tolist()]
        nulls_any = [bool(x) for x in has_nulls.cpu().tolist()]
        n_out = sum(recv_counts)
        out_blocks = []
        for b, nullable in zip(part_page.blocks, nulls_any):
            k = 2 if b.type == N.LONG_DECIMAL else 1
            src = self._as_tensor(b.values).view(-1, k) if k == 2 else self._as_tensor(b.values)
            dst = torch.empty((n_out, k) if k == 2 else (n_out,), dtype=src.dtype, device=self.device)
            dist.all_to_all_single(dst, src[:sum(counts)], output_split_sizes=recv_counts, input_split_sizes=counts, group=self.group)
            self.bytes_sent += (sum(counts) - counts[self.rank]) * src.element_size() * k
            nulls = None
            if nullable:
                ns = self._as_tensor(b.nulls) if b.nulls is not None else torch.zeros(sum(counts), dtype=torch.uint8, device=self.device)
                nulls = torch.empty(n_out, dtype=torch.uint8, device=self.device)
                dist.all_to_all_single(nulls, ns[:sum(counts)], output_split_sizes=recv_counts, input_split_sizes=counts, group=self.group)
            vals = dst.reshape(-1)
            if self.device == "cpu":
                vals = vals.numpy()
                if b.type == N.LONG_DECIMAL:
                    vals = vals.view(np.uint64)
                nulls = None if nulls is None else nulls.numpy()
            out_blocks.append(Block(b.type, vals, nulls, None, n_out, b.precision, b.scale))
        self.sent_rows += sum(counts) - counts[self.rank]
        self.received_rows += n_out
        return Page(*out_blocks, position_count=n_out)
