"""The allocation-free LM-head/argmax entry point (llmc_lmhead_argmax_noalloc, include/llmc_ops.h).

It serves callers that may not allocate GPU memory themselves or through us -- a process whose device
memory is pre-carved by its own arena, a plugin living inside a framework that owns all allocations,
a real-time loop that bans cudaMalloc/cudaFree after start-up.  They pass device buffers they already
hold; the kernel's scratch lives in a fixed array inside the library.

Checked here through the C ABI: its logits and tokens equal those of the existing allocating entry
point (llmc_op_lmhead) on two calls in a row (the second call finds the scratch as the first left it)
and, directly, a float64 numpy RMSNorm + matvec of the same inputs; sizes beyond the compiled-in
maximum are refused.  Device buffers come from torch (test plumbing
only).  This file sorts last so that a run which stops at the first failure has already run every
other test."""
import numpy as np
import pytest

from oracle import llama_np as lnp

pytestmark = pytest.mark.gpu


def _dev(t, a):
    return t.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("V,K,B", [(128256, 2048, 1), (1000, 512, 3), (4096, 256, 8)])
def test_noalloc_lmhead_matches_allocating_entry_point_twice(gpu, V, K, B):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(V + B)
    W = lnp.bf16_rne_bits((rng.standard_normal((V, K)) * 0.02).astype(np.float32))
    gain = lnp.bf16_rne_bits((1 + 0.1 * rng.standard_normal(K)).astype(np.float32))
    dW = _dev(torch, W.view(np.int16))
    dg = _dev(torch, gain.view(np.int16))
    for call in range(2):                                   # two calls in a row share the built-in scratch
        x = (rng.standard_normal((B, K)) * (2 + call)).astype(np.float32)
        want_tok, want_logits = gpu.op_lmhead(W, x, gain)   # the existing, allocating entry point
        dx = _dev(torch, x)
        dlog = torch.full((B, V), float("nan"), dtype=torch.float32, device="cuda")
        dtok = torch.full((B,), -7, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        rc = gpu.lib().llmc_lmhead_argmax_noalloc(0, V, K, B, dW.data_ptr(), dx.data_ptr(), dg.data_ptr(), 1e-5,
                                                  dlog.data_ptr(), dtok.data_ptr(), None)
        assert rc == 0, gpu.lib().llmc_noalloc_last_error()
        torch.cuda.synchronize()
        got = dlog.cpu().numpy()
        scale = float(np.sqrt(np.mean(want_logits.astype(np.float64) ** 2)))
        assert np.abs(got - want_logits).max() <= 1e-4 * scale      # the tolerance of the other decode-kernel tests
        # and against the arithmetic itself, not only against another CUDA entry point
        x64 = x.astype(np.float64)
        xn = x64 / np.sqrt((x64 * x64).mean(axis=1, keepdims=True) + 1e-5) * lnp.bf16_to_f32(gain).astype(np.float64)
        exact = xn @ lnp.bf16_to_f32(W).astype(np.float64).T
        assert np.abs(got - exact).max() <= 1e-4 * scale
        margin = np.sort(exact, axis=1)
        for b in range(B):                                          # argmax, wherever float64 has a clear winner
            if margin[b, -1] - margin[b, -2] > 1e-4 * scale:
                assert int(dtok.cpu().numpy()[b]) == int(np.argmax(exact[b]))
        assert dtok.cpu().numpy().tolist() == want_tok.tolist()
        # without a logits buffer the tokens are the same
        dtok2 = torch.full((B,), -7, dtype=torch.int32, device="cuda")
        rc = gpu.lib().llmc_lmhead_argmax_noalloc(0, V, K, B, dW.data_ptr(), dx.data_ptr(), dg.data_ptr(), 1e-5,
                                                  None, dtok2.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        assert dtok2.cpu().numpy().tolist() == want_tok.tolist()


def test_noalloc_refuses_sizes_beyond_its_scratch(gpu):
    torch = pytest.importorskip("torch")
    d = torch.zeros(16, dtype=torch.int32, device="cuda")
    p = d.data_ptr()
    lib = gpu.lib()
    assert lib.llmc_lmhead_argmax_noalloc(0, 262144 + 4, 256, 1, p, p, p, 1e-5, None, p, None) == gpu.ERR_LIMIT
    assert b"262144" in lib.llmc_noalloc_last_error()
    assert lib.llmc_lmhead_argmax_noalloc(0, 1024, 256, 9, p, p, p, 1e-5, None, p, None) == gpu.ERR_LIMIT
    assert lib.llmc_lmhead_argmax_noalloc(0, 1024, 256, 0, p, p, p, 1e-5, None, p, None) == gpu.ERR_LIMIT
    assert lib.llmc_lmhead_argmax_noalloc(0, 1024, 250, 1, p, p, p, 1e-5, None, p, None) == gpu.ERR_INVALID
    assert lib.llmc_lmhead_argmax_noalloc(0, 1024, 256, 1, None, p, p, 1e-5, None, p, None) == gpu.ERR_INVALID
