"""TF32 matmul peak with the protocol MEASURED_PEAKS.json uses for BF16: torch.matmul 8192^3,
allow_tf32, best of 10 (burst) and back to back for 4 s (sustained).  A denominator for TF32-mode fractions."""
import json, time, torch
torch.backends.cuda.matmul.allow_tf32 = True
n = 8192
a = torch.randn(n, n, device="cuda"); b = torch.randn(n, n, device="cuda")
for _ in range(3):
    a @ b
torch.cuda.synchronize()
best = 1e9
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
flop = 2 * n ** 3
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.time(); k = 0
e0.record()
while time.time() - t0 < 4.0:
    for _ in range(20):
        a @ b
    k += 20
    torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
out = {"tf32_tflops": flop / (best * 1e-3) / 1e12, "tf32_tflops_sustained": flop * k / (e0.elapsed_time(e1) * 1e-3) / 1e12,
       "how": "torch.matmul fp32 8192^3 with allow_tf32 (2*N^3): best of 10 (burst) and back to back for 4 s (sustained)"}
torch.backends.cuda.matmul.allow_tf32 = False
a16, b16 = a.bfloat16(), b.bfloat16()
for _ in range(3):
    a16 @ b16
torch.cuda.synchronize()
best = 1e9
for _ in range(10):
    e0.record(); a16 @ b16; e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
out["bf16_tflops_same_box"] = flop / (best * 1e-3) / 1e12
print(json.dumps(out))
