"""svb_asnorm_stats_static: the AS-norm top-K statistics with the workspace in the library's own file-scope device array
(no allocation inside the call; include/svb.h).  No counterpart in the reference.
Collected last on purpose."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _stats_pair(nv, emb, coh, top):
    """(mu, sigma) through the existing entry point with a caller workspace, and through the static-scratch variant."""
    U, D = emb.shape
    N = coh.shape[0]
    need = int(nv.lib().svb_asnorm_workspace_bytes(U, N, D))
    assert need <= int(nv.lib().svb_asnorm_static_bytes())
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    ref = torch.empty((2, U), dtype=torch.float32, device="cuda")
    got = torch.full((2, U), float("nan"), dtype=torch.float32, device="cuda")
    nv.check(nv.lib().svb_asnorm_stats(nv.ptr(emb), U, nv.ptr(coh), N, D, top, nv.ptr(ref[0]), nv.ptr(ref[1]), nv.ptr(ws),
                                       need, nv.stream_ptr()))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    nv.check(nv.lib().svb_asnorm_stats_static(nv.ptr(emb), U, nv.ptr(coh), N, D, top, nv.ptr(got[0]), nv.ptr(got[1]),
                                              nv.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
    return ref, got


def test_static_scratch_matches_workspace_entry_point_twice():
    from speakerverification_b200 import _native as nv, synth
    assert int(nv.lib().svb_asnorm_static_bytes()) == 16 << 20
    # two calls in a row on the shared array: the second one is SMALLER and of another shape, so it runs on top of
    # the first call's leftovers (operand padding, score chunk, counters)
    for U, N, D, top, seed in ((1000, 2000, 192, 300, 7), (333, 1001, 100, 50, 8), (1000, 2000, 192, 300, 9)):
        e, _ = synth.scoring_embeddings(U, 40, dim=D, seed=seed)
        c, _ = synth.scoring_embeddings(N, 200, dim=D, seed=seed + 100)
        emb, coh = torch.from_numpy(e).cuda(), torch.from_numpy(c).cuda()
        ref, got = _stats_pair(nv, emb, coh, top)
        # same kernels on the same inputs: equal to the workspace entry point (the tests' kernel-exactness bound is 2e-6)
        assert (ref - got).abs().max().item() <= 2e-6, (U, N, D, (ref - got).abs().max().item())
        S = -np.sort(-(e.astype(np.float64) @ c.astype(np.float64).T), axis=1)[:, :top]
        assert np.abs(got[0].cpu().numpy() - S.mean(1)).max() <= 2e-6
        assert np.abs(got[1].cpu().numpy() - S.std(1)).max() <= 2e-6


def test_static_scratch_rejects_sizes_beyond_the_array():
    from speakerverification_b200 import _native as nv
    U, N, D = 4096, 20000, 192
    assert int(nv.lib().svb_asnorm_workspace_bytes(U, N, D)) > int(nv.lib().svb_asnorm_static_bytes())
    emb = torch.zeros((U, D), device="cuda")
    coh = torch.zeros((N, D), device="cuda")
    out = torch.zeros((2, U), device="cuda")
    rc = nv.lib().svb_asnorm_stats_static(nv.ptr(emb), U, nv.ptr(coh), N, D, 300, nv.ptr(out[0]), nv.ptr(out[1]),
                                          nv.stream_ptr())
    assert rc != 0
    msg = nv.lib().svb_last_error()
    msg = msg.decode() if isinstance(msg, (bytes, bytearray)) else str(msg)
    assert "svb_asnorm_stats_static" in msg and "16777216" in msg
