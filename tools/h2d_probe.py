import torch, time
n = 4 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2):
    torch.cuda.synchronize(); t = time.perf_counter(); d.copy_(h, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print("H2D pinned GB/s", n / dt / 1e9)
hp = torch.empty(n, dtype=torch.uint8)
torch.cuda.synchronize(); t = time.perf_counter(); d.copy_(hp); torch.cuda.synchronize(); print("H2D pageable GB/s", n / (time.perf_counter() - t) / 1e9)
t = time.perf_counter(); h2 = d.cpu(); print("D2H pageable GB/s", n / (time.perf_counter() - t) / 1e9)
