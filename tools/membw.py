import torch
n = 1 << 30
a = torch.empty(n, dtype=torch.float32, device="cuda")
b = torch.empty(n, dtype=torch.float32, device="cuda")
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms = t(lambda: a.fill_(1.0)); print("fill_  4GiB write-only: %.3f ms  %.0f GB/s" % (ms, 4 * n / ms / 1e6))
ms = t(lambda: a.zero_()); print("zero_  4GiB (memset)  : %.3f ms  %.0f GB/s" % (ms, 4 * n / ms / 1e6))
ms = t(lambda: b.copy_(a)); print("copy_  4GiB r+w       : %.3f ms  %.0f GB/s (read+write)" % (ms, 8 * n / ms / 1e6))
ms = t(lambda: a.sum()); print("sum    4GiB read-only : %.3f ms  %.0f GB/s" % (ms, 4 * n / ms / 1e6))
