"""Helper run in a SUBPROCESS by tests/test_kernels_gpu.py so that a hung tensor-core kernel
(an mbarrier that never completes) cannot hang the whole test run.  Prints one JSON line."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc  # noqa: E402
from oracle import llama_np as lnp  # noqa: E402


def main():
    T, N, K, epi, impl = map(int, sys.argv[1:6])
    split = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    rng = np.random.default_rng(T * 7 + N * 3 + K + epi)
    a32 = rng.standard_normal((T, K)).astype(np.float32)
    A = lnp.bf16_rne_bits(a32)
    Alo = lnp.bf16_rne_bits(a32 - lnp.bf16_to_f32(A)) if split else None    # a32 ~= A + Alo (16 mantissa bits)
    W = lnp.bf16_rne_bits((rng.standard_normal((N, K)) * 0.05).astype(np.float32))
    Af, Wf = lnp.bf16_to_f32(A).astype(np.float64), lnp.bf16_to_f32(W).astype(np.float64)
    if split:
        Af = Af + lnp.bf16_to_f32(Alo).astype(np.float64)
    ref = Af @ Wf.T
    res = {"T": T, "N": N, "K": K, "epi": epi, "impl": impl, "split": split}
    if epi == 2:
        base = rng.standard_normal((T, N)).astype(np.float32)
        out = llmc.op_gemm(A, W, epi=2, impl=impl, out_in=base, A_lo=Alo)
        ref = ref + base
        got = out.astype(np.float64)
    elif epi == 3:
        I = N // 2
        g, u = ref[:, :I], ref[:, I:]
        ref = g / (1.0 + np.exp(-g)) * u
        hi, lo = llmc.op_gemm(A, W, epi=3, impl=impl, A_lo=Alo, want_lo=True)
        got = lnp.bf16_to_f32(hi).astype(np.float64)
        both = got + lnp.bf16_to_f32(lo).astype(np.float64)
        res["pair_max_excess"] = float((np.abs(both - ref) - (4e-4 * np.sqrt((ref ** 2).mean()) + 1e-5
                                                              + 2.0 ** -15 * np.abs(ref))).max())
    elif epi == 0:
        got = lnp.bf16_to_f32(llmc.op_gemm(A, W, epi=0, impl=impl, A_lo=Alo)).astype(np.float64)
    else:
        got = llmc.op_gemm(A, W, epi=1, impl=impl, A_lo=Alo).astype(np.float64)
    err = np.abs(got - ref)
    res["max_abs"] = float(err.max())
    res["ref_rms"] = float(np.sqrt((ref ** 2).mean()))
    # per-element tolerance: fp32 accumulation noise, plus one bf16 rounding (2^-9 |value|, with
    # margin 2^-8) where the kernel's OUTPUT is bf16
    tol = 4e-4 * res["ref_rms"] + 1e-5 + (2.0 ** -8 * np.abs(ref) if epi in (0, 3) else 0.0)
    res["max_excess"] = float((err - tol).max())
    res["bad_rows"] = int((err.max(axis=1) > 0.05 * res["ref_rms"] + 1e-3).sum())
    res["first_bad_cols"] = np.nonzero(err.max(axis=0) > 0.05 * res["ref_rms"] + 1e-3)[0][:8].tolist()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
