"""HBM write-only / read-only / copy bandwidth with torch (context for the training forward's stash stores)."""
import torch
x = torch.empty(8 << 30, dtype=torch.uint8, device="cuda")
y = torch.empty(8 << 30, dtype=torch.uint8, device="cuda")
def t(f, n=5):
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(n):
        a.record(); f(); b.record(); torch.cuda.synchronize(); best = min(best, a.elapsed_time(b))
    return best
ms = t(lambda: x.fill_(1)); print("write-only fill_  : %.0f GB/s" % (x.numel() / ms / 1e6))
ms = t(lambda: torch.cuda.memset(x.data_ptr(), 0, x.numel()) if hasattr(torch.cuda, "memset") else x.zero_()); print("write-only zero_  : %.0f GB/s" % (x.numel() / ms / 1e6))
xv = x.view(torch.float32)
ms = t(lambda: xv.sum()); print("read-only sum     : %.0f GB/s" % (x.numel() / ms / 1e6))
ms = t(lambda: y.copy_(x)); print("copy (read+write) : %.0f GB/s" % (2 * x.numel() / ms / 1e6))
