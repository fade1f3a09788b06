"""Kernel-level parity (GPU): every kernel of the hot path, driven through the C ABI hooks of
include/llmc_ops.h, against the numpy oracle pieces (oracle/llama_np.py).

Tolerances: the decode kernels compute in fp32 from bf16 weights exactly like the oracle, so only
summation order differs (rel 1e-4 of the output scale).  Kernels whose OUTPUT is bf16 add one
bf16 rounding (2^-8 relative)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import llama_np as lnp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rnd_bf16(rng, shape, scale=1.0):
    return lnp.bf16_rne_bits((rng.standard_normal(shape) * scale).astype(np.float32))


def close(got, want, rel, what=""):
    scale = float(np.sqrt(np.mean(np.asarray(want, np.float64) ** 2))) + 1e-30
    err = float(np.abs(np.asarray(got, np.float64) - np.asarray(want, np.float64)).max())
    assert err <= rel * scale, "%s: max-abs %.3e > %.1e * rms %.3e" % (what, err, rel, scale)


@pytest.mark.parametrize("N,K,B", [(512, 256, 1), (2048, 2048, 1), (3072, 2048, 1), (2048, 8192, 1), (1000, 512, 3),
                                   (768, 1024, 2), (2048, 2048, 4), (512, 4096, 8), (6, 256, 5)])
@pytest.mark.parametrize("norm,resid", [(0, 0), (1, 0), (0, 1)])
def test_gemv(gpu, N, K, B, norm, resid):
    rng = np.random.default_rng(N + K + B)
    W = rnd_bf16(rng, (N, K), 0.02)
    x = (rng.standard_normal((B, K)) * 1.7).astype(np.float32)
    gain = lnp.bf16_rne_bits((1 + 0.1 * rng.standard_normal(K)).astype(np.float32))
    y0 = rng.standard_normal((B, N)).astype(np.float32)
    xh = lnp.rmsnorm(x, lnp.bf16_to_f32(gain)) if norm else x
    want = xh.astype(np.float64) @ lnp.bf16_to_f32(W).astype(np.float64).T + (y0 if resid else 0)
    got = gpu.op_gemv(W, x, gain if norm else None, y_in=y0 if resid else None)
    close(got, want, 1e-4, "gemv")


@pytest.mark.parametrize("I,K,B", [(512, 256, 1), (8192, 2048, 1), (768, 512, 3), (1024, 1024, 8)])
def test_gemv_swiglu(gpu, I, K, B):
    rng = np.random.default_rng(I + K + B)
    Wg, Wu = rnd_bf16(rng, (I, K), 0.03), rnd_bf16(rng, (I, K), 0.03)
    x = (rng.standard_normal((B, K)) * 2).astype(np.float32)
    gain = lnp.bf16_rne_bits((1 + 0.1 * rng.standard_normal(K)).astype(np.float32))
    xh = lnp.rmsnorm(x, lnp.bf16_to_f32(gain)).astype(np.float64)
    g = xh @ lnp.bf16_to_f32(Wg).astype(np.float64).T
    u = xh @ lnp.bf16_to_f32(Wu).astype(np.float64).T
    close(gpu.op_gemv_swiglu(Wg, Wu, x, gain), g / (1 + np.exp(-g)) * u, 1e-4, "swiglu")


@pytest.mark.parametrize("V,K,B", [(1024, 256, 1), (128256, 2048, 1), (1000, 512, 3), (2048, 512, 8)])
def test_lmhead_logits_and_argmax(gpu, V, K, B):
    rng = np.random.default_rng(V + K + B)
    W = rnd_bf16(rng, (V, K), 0.02)
    x = (rng.standard_normal((B, K)) * 3).astype(np.float32)
    gain = lnp.bf16_rne_bits((1 + 0.1 * rng.standard_normal(K)).astype(np.float32))
    want = lnp.rmsnorm(x, lnp.bf16_to_f32(gain)).astype(np.float64) @ lnp.bf16_to_f32(W).astype(np.float64).T
    tok, lg = gpu.op_lmhead(W, x, gain)
    close(lg, want, 1e-4, "lm head")
    assert tok.tolist() == [int(np.argmax(r)) for r in lg]       # argmax of the GPU's own logits, exactly
    tok2, _ = gpu.op_lmhead(W, x, gain, want_logits=False)
    assert tok2.tolist() == tok.tolist()


def test_lmhead_ties_resolve_to_lowest_index(gpu):
    V, K = 4096, 256
    W = np.zeros((V, K), np.uint16)
    row = lnp.bf16_rne_bits(np.full(K, 0.5, np.float32))
    for r in (3000, 17, 2049, 4095):                               # identical rows -> identical logits
        W[r] = row
    x = np.ones((2, K), np.float32)
    gain = lnp.bf16_rne_bits(np.ones(K, np.float32))
    tok, lg = gpu.op_lmhead(W, x, gain)
    assert tok.tolist() == [17, 17]
    assert lg[0, 17] == lg[0, 3000] == lg[0, 4095] > 0


@pytest.mark.parametrize("nq,nkv,d", [(32, 8, 64), (32, 8, 128), (8, 2, 32), (4, 4, 64), (16, 2, 64), (4, 2, 128)])
@pytest.mark.parametrize("B,ctxs,page,nsplit", [(1, [1], 64, 0), (1, [33], 32, 0), (1, [640], 64, 0), (1, [700], 64, 5),
                                                  (3, [5, 129, 64], 64, 0), (2, [257, 31], 128, 1), (1, [2500], 64, 0),
                                                  (2, [3000, 10], 64, 8)])
@pytest.mark.parametrize("fp32", [True, False])
def test_attn_decode(gpu, nq, nkv, d, B, ctxs, page, nsplit, fp32):
    rng = np.random.default_rng(nq * d + sum(ctxs))
    max_ctx = ((max(ctxs) + page - 1) // page) * page
    q = rng.standard_normal((B, nq * d)).astype(np.float32)
    if fp32:
        k = rng.standard_normal((B, max_ctx, nkv * d)).astype(np.float32)
        v = rng.standard_normal((B, max_ctx, nkv * d)).astype(np.float32)
        kf, vf = k, v
    else:
        k = rnd_bf16(rng, (B, max_ctx, nkv * d), 1.0)
        v = rnd_bf16(rng, (B, max_ctx, nkv * d), 1.0)
        kf, vf = lnp.bf16_to_f32(k), lnp.bf16_to_f32(v)
    got = gpu.op_attn_decode(q, k, v, ctxs, nq, nkv, d, page, nsplit)
    assert np.isfinite(got).all()
    for b in range(B):
        kk = kf[b, :ctxs[b]].reshape(ctxs[b], nkv, d)
        vv = vf[b, :ctxs[b]].reshape(ctxs[b], nkv, d)
        want = lnp.attention(q[b].reshape(nq, d), kk, vv).reshape(-1)
        close(got[b], want, 2e-4, "attn decode b=%d" % b)


@pytest.mark.parametrize("T,H", [(1, 256), (37, 2048), (130, 4096)])
def test_rmsnorm_rows(gpu, T, H):
    rng = np.random.default_rng(T + H)
    x = (rng.standard_normal((T, H)) * 2.5).astype(np.float32)
    gain = lnp.bf16_rne_bits((1 + 0.1 * rng.standard_normal(H)).astype(np.float32))
    got = lnp.bf16_to_f32(gpu.op_rmsnorm_rows(x, gain))
    want = lnp.rmsnorm(x, lnp.bf16_to_f32(gain))
    assert np.abs(got - want).max() <= 2 ** -8 * np.abs(want).max() + 1e-6
    hi, lo = gpu.op_rmsnorm_rows(x, gain, split=True)                  # hi + lo carries 16 mantissa bits
    assert np.array_equal(hi, gpu.op_rmsnorm_rows(x, gain))
    pair = lnp.bf16_to_f32(hi) + lnp.bf16_to_f32(lo)
    assert np.abs(pair - want).max() <= 2 ** -15 * np.abs(want).max() + 1e-6


@pytest.mark.parametrize("impl", [0, 1, 2])   # 0: tcgen05 kernel where the shape allows it (else mma.sync); 1: CUDA-core kernel; 2: mma.sync kernel
@pytest.mark.parametrize("nq,nkv,d", [(8, 2, 32), (32, 8, 64), (4, 2, 128), (4, 4, 64), (16, 4, 128), (8, 1, 64)])
@pytest.mark.parametrize("T,p0,page", [(1, 0, 64), (20, 0, 32), (70, 0, 64), (9, 100, 64), (130, 61, 32),
                                       (130, 61, 64), (33, 200, 128)])
@pytest.mark.parametrize("fp32", [True, False])
def test_rope_append_and_prefill_attention(gpu, nq, nkv, d, T, p0, page, fp32, impl):
    rng = np.random.default_rng(nq + d + T + p0)
    QD, KD = nq * d, nkv * d
    qkv = rng.standard_normal((T, QD + 2 * KD)).astype(np.float32)
    if fp32:
        kp = rng.standard_normal((p0, KD)).astype(np.float32) if p0 else None
        vp = rng.standard_normal((p0, KD)).astype(np.float32) if p0 else None
        widen = lambda a: a
    else:
        kp = rnd_bf16(rng, (p0, KD)) if p0 else None
        vp = rnd_bf16(rng, (p0, KD)) if p0 else None
        widen = lnp.bf16_to_f32
    out, lo, qo, kn, vn = gpu.op_prefill_attention(qkv, kp, vp, nq, nkv, d, page, kv_fp32=fp32, split=True, impl=impl)
    cos, sin = lnp.rope_tables(p0 + T, d)
    pos = np.arange(p0, p0 + T)
    q_want = lnp.rope(qkv[:, :QD].reshape(T, nq, d), pos, cos, sin)
    k_want = lnp.rope(qkv[:, QD:QD + KD].reshape(T, nkv, d), pos, cos, sin)
    v_want = qkv[:, QD + KD:].reshape(T, nkv, d)
    store_tol = 1e-6 if fp32 else 2 ** -8
    assert np.abs(qo.reshape(T, nq, d) - q_want).max() <= 1e-6 * np.abs(q_want).max() + 1e-6
    assert np.abs(widen(kn).reshape(T, nkv, d) - k_want).max() <= store_tol * np.abs(k_want).max() + 1e-6
    assert np.abs(widen(vn).reshape(T, nkv, d) - v_want).max() <= store_tol * np.abs(v_want).max() + 1e-6
    # attention on exactly what the GPU stored (so only its own arithmetic is under test)
    k_all = widen(kn).reshape(T, nkv, d)
    v_all = widen(vn).reshape(T, nkv, d)
    if p0:
        k_all = np.concatenate([widen(kp).reshape(p0, nkv, d), k_all])
        v_all = np.concatenate([widen(vp).reshape(p0, nkv, d), v_all])
    qf = qo.reshape(T, nq, d)
    hi = lnp.bf16_to_f32(out).reshape(T, nq, d)
    pair = hi + lnp.bf16_to_f32(lo).reshape(T, nq, d)
    for t in range(T):
        want = lnp.attention(qf[t], k_all[:p0 + t + 1], v_all[:p0 + t + 1])
        assert np.abs(hi[t] - want).max() <= 2 ** -8 * np.abs(want).max() + 1e-4, t
        assert np.abs(pair[t] - want).max() <= 2e-4 * np.abs(want).max() + 1e-5, t


def test_prefill_attention_tensor_core_matches_cuda_core(gpu):
    """Full-size head geometry (32 q / 8 kv heads x 64, and 128), a prompt longer than several tiles, with a
    cached prefix: the tensor-core kernel (3 x tf32 split products) must agree with the CUDA-core fp32 kernel to
    the rounding of the bf16 hi+lo output pair."""
    for nq, nkv, d in [(32, 8, 64), (32, 8, 128)]:
        rng = np.random.default_rng(d)
        T, p0 = 300, 77
        QD, KD = nq * d, nkv * d
        qkv = rng.standard_normal((T, QD + 2 * KD)).astype(np.float32)
        kp = rng.standard_normal((p0, KD)).astype(np.float32)
        vp = rng.standard_normal((p0, KD)).astype(np.float32)
        res = [gpu.op_prefill_attention(qkv, kp, vp, nq, nkv, d, 64, kv_fp32=True, split=True, impl=i) for i in (0, 1, 2)]
        pair = [lnp.bf16_to_f32(r[0]) + lnp.bf16_to_f32(r[1]) for r in res]
        assert np.abs(pair[0] - pair[1]).max() <= 2e-5 * np.abs(pair[1]).max()     # tcgen05 kernel vs CUDA cores
        assert np.abs(pair[2] - pair[1]).max() <= 2e-5 * np.abs(pair[1]).max()     # mma.sync kernel vs CUDA cores


def run_gemm_check(T, N, K, epi, impl, split=0, timeout=120):
    cmd = [sys.executable, os.path.join(ROOT, "tests", "_gemm_tc_check.py")] + \
        [str(v) for v in (T, N, K, epi, impl, split)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, "gemm check failed:\n" + r.stdout[-2000:] + r.stderr[-4000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


GEMM_CASES = [(128, 384, 256, 1), (20, 256, 256, 1), (512, 3072, 2048, 1), (512, 2048, 2048, 2), (300, 1024, 512, 3),
              (512, 16384, 2048, 3), (512, 2048, 8192, 2), (130, 512, 1024, 0), (1024, 6144, 4096, 1)]


@pytest.mark.parametrize("split", [0, 1])
@pytest.mark.parametrize("T,N,K,epi", GEMM_CASES)
def test_gemm_simple_comparator(gpu, T, N, K, epi, split):
    if T * N * K > 2 ** 34:
        pytest.skip("large case only on the tensor-core kernel")
    r = run_gemm_check(T, N, K, epi, impl=1, split=split)
    assert r["bad_rows"] == 0, r
    assert r["max_excess"] <= 0, r
    assert r.get("pair_max_excess", 0) <= 0, r


# decode-shaped GEMMs (one M tile): impl 2 = tcgen05 with split-K allowed (partial sums added by fp32 reductions)
SPLITK_CASES = [(128, 2048, 8192, 2), (17, 2048, 2048, 2), (64, 3072, 2048, 1), (128, 1024, 512, 1), (9, 4096, 4096, 2),
                (100, 6144, 4096, 1)]


@pytest.mark.parametrize("split", [0, 1])
@pytest.mark.parametrize("T,N,K,epi", SPLITK_CASES)
def test_gemm_tcgen05_split_k(gpu, T, N, K, epi, split):
    r = run_gemm_check(T, N, K, epi, impl=2, split=split)
    assert r["bad_rows"] == 0, r
    assert r["max_excess"] <= 0, r


# the decode GEMM (impl 3: weights as the M tile, the batch as N = 16/32/64/128)
DECODE_GEMM_CASES = SPLITK_CASES + [(1, 256, 128, 1), (16, 16384, 2048, 3), (5, 1024, 512, 3), (33, 2048, 256, 3),
                                    (128, 4096, 2048, 3), (70, 128256, 2048, 1), (40, 512, 14336, 2)]


@pytest.mark.parametrize("split", [0, 1])
@pytest.mark.parametrize("T,N,K,epi", DECODE_GEMM_CASES)
def test_gemm_decode_tc(gpu, T, N, K, epi, split):
    r = run_gemm_check(T, N, K, epi, impl=3, split=split, timeout=300)
    assert r["bad_rows"] == 0, r
    assert r["max_excess"] <= 0, r
    assert r.get("pair_max_excess", 0) <= 0, r


@pytest.mark.parametrize("split", [0, 1])
@pytest.mark.parametrize("T,N,K,epi", GEMM_CASES)
def test_gemm_tcgen05(gpu, T, N, K, epi, split):
    """split=1: the A operand is a bf16 hi+lo pair (the default precision of the engine's prefill)."""
    r = run_gemm_check(T, N, K, epi, impl=0, split=split)
    assert r["bad_rows"] == 0, r
    assert r["max_excess"] <= 0, r
    assert r.get("pair_max_excess", 0) <= 0, r
