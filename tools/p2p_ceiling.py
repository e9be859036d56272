"""Under torchrun with 2+ GPUs: what a plain peer copy reaches over NVLink (the ceiling for the exchange's stores).
DMA = cudaMemcpyPeer through the copy engines; SM = an elementwise kernel storing into the peer's buffer (16-byte vector stores)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
n = 300_000_000          # int64 elements = 2.4 GB
buf = symm.empty(n, dtype=torch.int64, device="cuda")
hdl = symm.rendezvous(buf, dist.group.WORLD)
src = torch.arange(n, dtype=torch.int64, device="cuda")
peer = hdl.get_buffer((rank + 1) % world, (n,), torch.int64)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for name, fn in (("DMA copy_ to peer", lambda: peer.copy_(src)),
                 ("SM kernel torch.add(out=peer)", lambda: torch.add(src, 1, out=peer)),
                 ("SM kernel, half of the rows to the peer, half local", lambda: (torch.add(src[: n // 2], 1, out=peer[: n // 2]), torch.add(src[n // 2:], 1, out=buf[n // 2:]))),
                 ("local copy", lambda: buf.copy_(src))):
    ms = timed(fn)
    if rank == 0:
        print("%-55s %7.3f ms  %7.1f GB/s (payload %0.1f GB, every rank at once)" % (name, ms, n * 8 / ms / 1e6, n * 8 / 1e9), flush=True)
dist.destroy_process_group()
