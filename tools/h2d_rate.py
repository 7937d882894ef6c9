"""Host -> device copy rate from page-locked memory by transfer size (what bounds sba_create's upload phase)."""
import torch
for mb in (4, 20, 80, 120, 1024):
    n = mb << 20
    src = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(n, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        dst.copy_(src, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    print("%5d MB: %.3f ms  %.1f GB/s" % (mb, ms, n / ms / 1e6))

# the same copies when the link has carried nothing for ~30 ms (the solve before the next sba_create), compute busy or idle
import time
n = 20 << 20
src = torch.empty(n, dtype=torch.uint8, pin_memory=True)
src2 = torch.empty(4 * n, dtype=torch.uint8, pin_memory=True)
dst = torch.empty(4 * n, dtype=torch.uint8, device="cuda")
x = torch.randn(8192, 8192, device="cuda")
for label, busy in (("after 30 ms of kernels", True), ("after 30 ms idle", False)):
    for rep in range(3):
        torch.cuda.synchronize()
        if busy:
            t0 = time.perf_counter()
            while time.perf_counter() - t0 < 0.03:
                y = x @ x
            torch.cuda.synchronize()
        else:
            time.sleep(0.03)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record(); dst[:n].copy_(src, non_blocking=True)
        ev[1].record(); dst[:n].copy_(src, non_blocking=True)
        ev[2].record(); dst.copy_(src2, non_blocking=True)
        ev[3].record()
        torch.cuda.synchronize()
        print("%s: 20 MB %.3f ms, 20 MB %.3f ms, 80 MB %.3f ms" % (label, ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])))
