"""Dense matmul peaks measured the way MEASURED_PEAKS.json was (torch.matmul 8192^3, 2 N^3 flops, best of 10 and 4 s back to
back), for TF32 next to bf16 and fp16: the denominators DESIGN.md §3 compares the operand formats with."""
import json, time, torch
torch.backends.cuda.matmul.allow_tf32 = True
N = 8192
out = {}
for name, dt in (("bf16", torch.bfloat16), ("fp16", torch.float16), ("tf32", torch.float32)):
    a = torch.randn(N, N, device="cuda", dtype=dt); b = torch.randn(N, N, device="cuda", dtype=dt)
    for _ in range(3):
        a @ b
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time(); n = 0
    e0.record()
    while time.time() - t0 < 4.0:
        for _ in range(20):
            a @ b
        n += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    out[name] = {"burst_tflops": 2 * N ** 3 / (best * 1e-3) / 1e12, "sustained_tflops": 2 * N ** 3 * n / (e0.elapsed_time(e1) * 1e-3) / 1e12}
print(json.dumps(out))
