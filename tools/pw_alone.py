"""gemm_pw alone on the encoder's pointwise shapes, CUDA events: (a) the product kernel (GELU + BN epilogue, fp16 out,
cta_group::2 instance), (b) the same mainloop with a store-free epilogue (svb_debug_gemm_noepi: 1-SM instance, TMEM reads
and compares only), (c) torch.matmul fp16 (cuBLAS) on the same shape -- separates operand feed from epilogue cost."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speakerverification_b200 import _native as nv


def timeit(run, reps=10):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


M = 592 * 309
for (K, N) in ((1024, 1024), (3072, 3072)):
    a = torch.randn(M, K, device="cuda").to(torch.float16)
    w = torch.randn(N, K, 1, device="cuda") / K ** 0.5
    wp = nv.pack_conv_weight(w, planes=1, f16=True)
    z, o = torch.zeros(N, device="cuda"), torch.ones(N, device="cuda")
    out = torch.empty(M, N, device="cuda", dtype=torch.float16)
    ms = timeit(lambda: nv.tdnn_layer(a, wp, N, K, 1, 1, z, o, z, out, act=nv.ACT_GELU, out_kind=nv.OUT_F16, w_f16=True, x_f16=True))
    print("gemm_pw (GELU+BN, f16 out, 2-SM) M=%d K=%d N=%d: %.1f us, %.0f TFLOP/s" % (M, K, N, ms * 1e3, 2.0 * M * K * N / ms / 1e9))
    thr = torch.full((M,), float("inf"), device="cuda")
    cnt = torch.zeros(M, dtype=torch.int32, device="cuda")
    cand = torch.zeros(M, device="cuda")
    ms = timeit(lambda: nv.check(nv.lib().svb_debug_gemm_noepi(nv.ptr(a), nv.ptr(wp), M, N, K, nv.ptr(thr), nv.ptr(cnt), nv.ptr(cand), nv.stream_ptr())))
    print("  store-free epilogue (1-SM instance):            %.1f us, %.0f TFLOP/s" % (ms * 1e3, 2.0 * M * K * N / ms / 1e9))
    wt = w[:, :, 0].half().t().contiguous()
    ms = timeit(lambda: torch.matmul(a, wt, out=out))
    print("  torch.matmul fp16 (cuBLAS):                       %.1f us, %.0f TFLOP/s" % (ms * 1e3, 2.0 * M * K * N / ms / 1e9))
