import torch, time
for mb in (6.29, 16, 64, 256):
    n = int(mb * 1e6)
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(5): d.copy_(h, non_blocking=True)
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(50): d.copy_(h, non_blocking=True)
        e1.record(s); s.synchronize()
    print("%.2f MB copies: %.1f GB/s" % (mb, n * 50 / e0.elapsed_time(e1) / 1e6))
