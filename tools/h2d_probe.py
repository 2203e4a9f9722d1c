import torch, time
n = 8 * 1000 * 1000  # one 8 MB block in doubles... = 1e6 doubles
host = torch.empty(1000, 1000 * 1000 // 1000 * 1000, dtype=torch.float64).pin_memory()  # 1000 blocks x 8 MB = 8 GB
dev = torch.empty_like(host, device="cuda")
torch.cuda.synchronize()
for label, fn in (("1D 8GB single copy", lambda: dev.copy_(host, non_blocking=True)),
                  ("1D per-block copies", lambda: [dev[i].copy_(host[i], non_blocking=True) for i in range(1000)])):
    for _ in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(label, f"{host.numel()*8/dt/1e9:.1f} GB/s")
# 2D: 1000x1000 blocks into a pitched destination (row pitch 597540 doubles)
dst = torch.empty(1000, 597540, dtype=torch.float64, device="cuda")
blk = host[:597].view(597, 1000, 1000)
def two_d():
    for j in range(597):
        dst[:, j*1000:(j+1)*1000].copy_(blk[j], non_blocking=True)
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); two_d(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("2D block -> pitched", f"{597*8e6/dt/1e9:.1f} GB/s")
