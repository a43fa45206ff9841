"""Measured FP64 ceiling on this box: cuBLAS DGEMM (library baseline, not on the product path)."""
import json, torch
torch.backends.cuda.matmul.allow_tf32 = False
out = {}
for n in (4096, 8192):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    for _ in range(3):
        torch.matmul(a, b)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out[f"dgemm_{n}_tflops"] = 2 * n**3 / best * 1e-9
    # sustained, 2 s
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    reps = max(3, int(2000 / best))
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b)
    e1.record(); torch.cuda.synchronize()
    out[f"dgemm_{n}_tflops_sustained"] = 2 * n**3 * reps / e0.elapsed_time(e1) * 1e-9
# cusolver potrf baseline (library GPU baseline B3)
for n in (1024, 4096, 8192):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    k = a @ a.T + n * torch.eye(n, dtype=torch.float64, device="cuda")
    for _ in range(2):
        torch.linalg.cholesky(k)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); torch.linalg.cholesky(k); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out[f"cusolver_potrf_{n}_ms"] = best
    out[f"cusolver_potrf_{n}_tflops"] = n**3 / 3 / best * 1e-9
kb = torch.randn(256, 1024, 1024, dtype=torch.float64, device="cuda")
kb = kb @ kb.transpose(1, 2) + 1024 * torch.eye(1024, dtype=torch.float64, device="cuda")
for _ in range(2):
    torch.linalg.cholesky(kb)
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record(); torch.linalg.cholesky(kb); e1.record(); torch.cuda.synchronize()
out["torch_potrf_batched_256x1024_ms"] = e0.elapsed_time(e1)
print(json.dumps(out, indent=1))
