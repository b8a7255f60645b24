import torch, time
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); h.fill_(65)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for chunk in (n, 256 << 20, 64 << 20, 32 << 20, 8 << 20):
    for rep in range(3):
        torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for o in range(0, n, chunk): d[o:o + chunk].copy_(h[o:o + chunk], non_blocking=True)
        e1.record(); torch.cuda.synchronize()
    print("chunk %4d MiB: %.2f GB/s" % (chunk >> 20, n / e0.elapsed_time(e1) / 1e6))
