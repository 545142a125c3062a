"""cuBLAS bf16 throughput on the three GELU-GEMM shapes of one 592-crop launch (library reference, not the product path)."""
import torch
M = 592 * 309
for (K, N) in ((1024, 1024), (3072, 3072), (3072, 128)):
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        torch.matmul(a, w.t(), out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        torch.matmul(a, w.t(), out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("cuBLAS M=%d K=%d N=%d: %.1f us, %.0f TFLOP/s (executed, %d rows incl. halo)" % (M, K, N, ms * 1e3, 2.0 * M * K * N / ms / 1e9, M))
