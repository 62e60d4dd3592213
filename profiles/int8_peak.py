"""Measured dense int8 tensor peak on this GPU: cuBLASLt int8 GEMM (torch._int_mm), 8192^3, best of 10 and a 3 s loop."""
import json, time, torch
n = 8192
a = torch.randint(-8, 8, (n, n), dtype=torch.int8, device="cuda")
b = torch.randint(-8, 8, (n, n), dtype=torch.int8, device="cuda")
for _ in range(3):
    torch._int_mm(a, b)
torch.cuda.synchronize()
best = 1e9
for _ in range(10):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); torch._int_mm(a, b); e.record(); torch.cuda.synchronize()
    best = min(best, s.elapsed_time(e))
t0 = time.time(); it = 0
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
while time.time() - t0 < 3.0:
    for _ in range(20):
        torch._int_mm(a, b)
    it += 20
    torch.cuda.synchronize()
e.record(); torch.cuda.synchronize()
ops = 2.0 * n ** 3
print(json.dumps({"int8_tops_burst": ops / (best * 1e-3) / 1e12, "int8_tops_sustained": ops * it / (s.elapsed_time(e) * 1e-3) / 1e12,
                  "how": "torch._int_mm (cuBLASLt) int8 8192^3, int32 accumulate"}))
