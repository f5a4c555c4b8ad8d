import torch, time
x = torch.empty(265420800, dtype=torch.uint8).pin_memory()
y = torch.empty_like(x, device='cuda')
for _ in range(3): y.copy_(x, non_blocking=True)
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): y.copy_(x, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print("H2D GB/s", 10*x.numel()/ (e0.elapsed_time(e1)*1e-3)/1e9)
xs=[torch.empty(8294400, dtype=torch.uint8).pin_memory() for _ in range(32)]
ys=[torch.empty(8294400, dtype=torch.uint8, device='cuda') for _ in range(32)]
e0.record()
for _ in range(10):
    for a,b in zip(xs,ys): b.copy_(a, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print("H2D per-frame copies GB/s", 10*32*8294400/ (e0.elapsed_time(e1)*1e-3)/1e9)
