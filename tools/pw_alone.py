"""gemm_pw alone on the same shapes (GELU + BN epilogue, bf16 out), CUDA events."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speakerverification_b200 import _native as nv
M = 592 * 309
for (K, N) in ((1024, 1024), (3072, 3072)):
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = torch.randn(N, K, 1, device="cuda") / K ** 0.5
    wp = nv.pack_conv_weight(w, planes=1)
    z, o = torch.zeros(N, device="cuda"), torch.ones(N, device="cuda")
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    run = lambda: nv.tdnn_layer(a, wp, N, K, 1, 1, z, o, z, out, act=nv.ACT_GELU)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("gemm_pw M=%d K=%d N=%d: %.1f us, %.0f TFLOP/s (executed)" % (M, K, N, ms * 1e3, 2.0 * M * K * N / ms / 1e9))
