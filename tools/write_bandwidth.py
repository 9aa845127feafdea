import torch, json
n = 2412610692
x = torch.empty(n, dtype=torch.float64, device="cuda")
res = {}
for name, fn in (("memset (cudaMemsetAsync via zero_)", lambda: x.zero_()), ("fill_(1.5) (torch vectorized store kernel)", lambda: x.fill_(1.5))):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10): fn()
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 10
    res[name] = {"ms": ms, "GBps": n * 8 / ms / 1e6}
y = torch.empty(n // 2, dtype=torch.float64, device="cuda"); z = torch.empty_like(y)
for _ in range(3): z.copy_(y)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(10): z.copy_(y)
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / 10
res["copy (read + write, 2 x 9.65 GB)"] = {"ms": ms, "GBps": n * 8 / ms / 1e6}
print(json.dumps(res))
