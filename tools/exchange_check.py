"""Run under torchrun with N GPUs: the peer-memory exchange must deliver exactly the rows (and order) the NCCL all-to-all path does.
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/exchange_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import presto_b200 as P

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
ctx = P.GpuContext(torch.cuda.current_device())
N = P.native
rng = np.random.default_rng(100 + rank)
ok = True
for trial, n in enumerate([0 if rank == 0 else 5, 1000, 300_000 + 17 * rank, 2_000_000]):
    keys = rng.integers(-10**12, 10**12, n).astype(np.int64)
    dec = rng.integers(-10**15, 10**15, n).astype(np.int64)
    dnull = (rng.random(n) < 0.1).astype(np.uint8) if (trial + rank) % 2 == 0 else None      # nulls on some ranks only
    dbl = rng.random(n)
    big = [int(x) * 10**20 for x in rng.integers(1, 10**9, n)]
    i32 = rng.integers(-2**31, 2**31 - 1, n).astype(np.int32)
    page = P.Page(P.Block.long_array(keys), P.Block.long_array(dec, dnull, P.SHORT_DECIMAL), P.Block.double_array(dbl),
                  P.Block.int128_array(big), P.Block.int_array(i32) if hasattr(P.Block, "int_array") else P.Block.long_array(i32.astype(np.int64)),
                  position_count=n).to_device()
    a = P.HashPartitionExchange(P.DevicePartitioner(ctx), peer="auto")
    b = P.HashPartitionExchange(P.DevicePartitioner(ctx), peer=False)
    if trial == 0 and rank == 0:
        print("peer mode:", a.peer is not None, "-", a.peer_why, flush=True)
    for rep in range(2):       # twice: the second call reuses the symmetric buffers
        pa = a.exchange(page, [0])
        pb = b.exchange(page, [0])
        same = pa.position_count == pb.position_count
        for x, y in zip(pa.blocks, pb.blocks):
            xv = x.values if torch.is_tensor(x.values) else torch.as_tensor(x.values)
            yv = y.values if torch.is_tensor(y.values) else torch.as_tensor(y.values)
            same = same and torch.equal(xv.view(torch.uint8).cpu(), yv.view(torch.uint8).cpu())
            same = same and ((x.nulls is None) == (y.nulls is None))
            if x.nulls is not None and y.nulls is not None:
                same = same and torch.equal(torch.as_tensor(x.nulls).cpu(), torch.as_tensor(y.nulls).cpu())
        ok = ok and same
        if not same:
            print("rank %d trial %d rep %d MISMATCH (%d vs %d rows)" % (rank, trial, rep, pa.position_count, pb.position_count), flush=True)
# ---- a page returned by exchange N must survive exchange N+1 (same slot): lease=False hands out its own memory; with lease=True the
# page is a view of the slot's receive buffers and a different slot must be used to keep it
def same_page(x, y):
    return x.position_count == y.position_count and all(
        torch.equal(torch.as_tensor(p.values).view(torch.uint8).cpu(), torch.as_tensor(q.values).view(torch.uint8).cpu()) for p, q in zip(x.blocks, y.blocks))


def make(seed, n):
    r = np.random.default_rng(seed)
    return P.Page(P.Block.long_array(r.integers(0, 10**9, n).astype(np.int64)), P.Block.long_array(r.integers(-10**9, 10**9, n).astype(np.int64)),
                  position_count=n).to_device()


p1, p2 = make(1000 + rank, 400_000), make(2000 + rank, 350_000 + rank)
ex = P.HashPartitionExchange(P.DevicePartitioner(ctx), peer="auto")
ref = P.HashPartitionExchange(P.DevicePartitioner(ctx), peer=False)
want1, want2 = ref.exchange(p1, [0]), ref.exchange(p2, [0])
held = ex.exchange(p1, [0])                        # default: a copy
nxt = ex.exchange(p2, [0])
torch.cuda.synchronize()
hold_ok = same_page(held, want1) and same_page(nxt, want2)
leased = ex.exchange(p1, [0], slot="a", lease=True)
other = ex.exchange(p2, [0], slot="b", lease=True)  # another slot: `leased` stays valid
torch.cuda.synchronize()
hold_ok = hold_ok and same_page(leased, want1) and same_page(other, want2)
if not hold_ok:
    print("rank %d: a held page changed under the next exchange" % rank, flush=True)
ok = ok and hold_ok
t = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("exchange_check:", "OK" if int(t.item()) == 1 else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 1 else 1)
