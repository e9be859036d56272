import torch, time
x = torch.empty(6_000_000_000, dtype=torch.uint8, device="cuda")
y = torch.empty(750_000_000, dtype=torch.int64, device="cuda")
def t(f, reps=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms = t(lambda: x.zero_()); print("memset 6 GB: %.3f ms  %.2f TB/s" % (ms, 6e9 / ms / 1e9))
ms = t(lambda: x.fill_(255)); print("fill_ u8 6 GB: %.3f ms  %.2f TB/s" % (ms, 6e9 / ms / 1e9))
ms = t(lambda: y.fill_(-1)); print("fill_ i64 6 GB: %.3f ms  %.2f TB/s" % (ms, 6e9 / ms / 1e9))
z = torch.empty_like(y)
ms = t(lambda: z.copy_(y)); print("copy 6+6 GB: %.3f ms  %.2f TB/s" % (ms, 12e9 / ms / 1e9))
