import torch
x = torch.empty(1<<28, dtype=torch.float32, device='cuda')  # 1 GiB
y = torch.empty(1<<28, dtype=torch.float32, device='cuda')
def t(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
ms = t(lambda: x.zero_()); print("memset 1 GiB: %.3f ms  %.0f GB/s" % (ms, 2**30 / ms / 1e6))
ms = t(lambda: x.fill_(1.5)); print("fill   1 GiB: %.3f ms  %.0f GB/s" % (ms, 2**30 / ms / 1e6))
ms = t(lambda: y.copy_(x)); print("copy   1 GiB: %.3f ms  %.0f GB/s (r+w)" % (ms, 2 * 2**30 / ms / 1e6))
ms = t(lambda: x.sum()); print("read   1 GiB: %.3f ms  %.0f GB/s" % (ms, 2**30 / ms / 1e6))
