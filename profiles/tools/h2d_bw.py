import torch, time
x = torch.empty(32 * 1024 * 1024, dtype=torch.uint8).pin_memory()
d = torch.empty_like(x, device="cuda")
for blocks in (1, 8, 32):
    n = x.numel() // blocks
    torch.cuda.synchronize()
    for rep in range(3):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for b in range(blocks):
            d[b * n:(b + 1) * n].copy_(x[b * n:(b + 1) * n], non_blocking=True)
        e.record(); torch.cuda.synchronize()
    print(blocks, "blocks: %.3f ms  %.1f GB/s" % (s.elapsed_time(e), x.numel() / s.elapsed_time(e) / 1e6))
