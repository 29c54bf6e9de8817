import torch, time
n = 1 << 30
d = torch.empty(n, dtype=torch.uint8, device='cuda')
h = torch.empty(n, dtype=torch.uint8).pin_memory()
for _ in range(2): h.copy_(d); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
best = 1e9
for _ in range(5):
    e0.record(); h.copy_(d, non_blocking=True); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print("D2H pinned GB/s", n / best / 1e6)
best = 1e9
for _ in range(5):
    e0.record(); d.copy_(h, non_blocking=True); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print("H2D pinned GB/s", n / best / 1e6)
