"""Generates tests/golden/oracle_1b_wide.npz: fp32 CPU oracle runs (oracle/llama_ref.cpp, Llama-3.2-1B shape, seed 0) for
a few sequences of the two wide-batch parity cases of tests/test_engine_gpu.py (B = 32 and B = 128 at contexts of
512-560 tokens, BASELINE.json configs[4]'s 1B points).  Each checked sequence costs the oracle about a minute, the
GPU test replays them in seconds.

For batch B the prompts are rng(1000 + B): lengths uniform in [512, 560), ids uniform in [0, V); sequences
CHECK[B] are run through the oracle for N_NEW + 1 greedy tokens.  Stored per checked sequence: the tokens, and for the
LAST step the top-16 logits with indices plus the logits at 512 fixed vocabulary positions (rng seed 7).

    python tests/golden/gen_oracle_1b_wide_fixture.py        # ~10 min on 8 cores
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref

V, N_NEW = 128256, 6
CHECK = {32: [0, 13, 31], 128: [0, 64, 127]}


def prompts_for(B):
    rng = np.random.default_rng(1000 + B)
    return [rng.integers(0, V, int(n)).astype(np.int32) for n in rng.integers(512, 560, B)]


if __name__ == "__main__":
    t0 = time.time()
    o = ref.Oracle("llama-1b", seed=0, max_ctx=640)
    probe = np.sort(np.random.default_rng(7).choice(V, 512, replace=False)).astype(np.int32)
    out = {"probe_idx": probe, "n_new": np.int32(N_NEW)}
    for B, idx in CHECK.items():
        ps = prompts_for(B)
        for i in idx:
            o.reset()
            toks, lg = o.generate(ps[i], N_NEW + 1, want_logits=True)
            last = lg[N_NEW]
            top = np.argsort(-last)[:16].astype(np.int32)
            margins = np.sort(lg, axis=1)
            out["b%d_s%d_tokens" % (B, i)] = toks.astype(np.int32)
            out["b%d_s%d_margins" % (B, i)] = (margins[:, -1] - margins[:, -2]).astype(np.float32)
            out["b%d_s%d_probe" % (B, i)] = last[probe].astype(np.float32)
            out["b%d_s%d_top_idx" % (B, i)] = top
            out["b%d_s%d_top_val" % (B, i)] = last[top].astype(np.float32)
            print("B=%d sequence %d (%d tokens): %s, min margin %.3e, %.0f s" %
                  (B, i, len(ps[i]), toks.tolist(), float((margins[:, -1] - margins[:, -2]).min()), time.time() - t0), flush=True)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_1b_wide.npz")
    np.savez_compressed(path, **out)
    print("wrote", path)
