import torch, time
x = torch.empty(256*1024*1024, dtype=torch.uint8).pin_memory()
y = torch.empty_like(x, device='cuda')
for n in (16*1024*1024, 256*1024*1024):
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4): y[:n].copy_(x[:n], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(n, 4*n/e0.elapsed_time(e1)/1e6, "GB/s H2D")
    e0.record()
    for _ in range(4): x[:n].copy_(y[:n], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(n, 4*n/e0.elapsed_time(e1)/1e6, "GB/s D2H")
