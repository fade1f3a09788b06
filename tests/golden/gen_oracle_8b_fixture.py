"""Generates tests/golden/oracle_8b_2048.npz: the fp32 CPU oracle (oracle/llama_ref.cpp) on BASELINE.json configs[3]'s
model -- Llama-3-8B shape, seed 0 -- for a 2 048-token prompt followed by 64 greedy tokens.  The full run costs tens of
CPU-minutes, so it is done once, here, and the GPU test (tests/test_engine_gpu.py::test_llama8b_2048_prompt_fixture)
replays it teacher-forced in seconds.

Stored per step (the prefill's last position is step 0): the greedy token, the top-16 logits with their indices, and the
logits at 512 fixed vocabulary positions (rng seed 7) -- 528 checked values per step instead of 128 256, small enough
to commit.

    python tests/golden/gen_oracle_8b_fixture.py        # ~20-40 min on 8 cores, ~18 GB of RAM
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref

N_PROMPT, N_NEW, V = 2048, 64, 128256
t0 = time.time()
o = ref.Oracle("llama-8b", seed=0, max_ctx=N_PROMPT + N_NEW + 8)
prompt = np.random.default_rng(3).integers(0, V, N_PROMPT).astype(np.int32)
toks, logits = o.generate(prompt, N_NEW, want_logits=True)
probe = np.sort(np.random.default_rng(7).choice(V, 512, replace=False)).astype(np.int32)
top_idx = np.argsort(-logits, axis=1)[:, :16].astype(np.int32)
top_val = np.take_along_axis(logits, top_idx, axis=1).astype(np.float32)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_8b_2048.npz")
np.savez_compressed(out, prompt_seed=np.int32(3), n_prompt=np.int32(N_PROMPT), tokens=toks.astype(np.int32),
                    probe_idx=probe, probe_val=logits[:, probe].astype(np.float32), top_idx=top_idx, top_val=top_val,
                    logit_std=np.float32(logits.std()))
print("wrote %s in %.0f s; tokens %s ..., logit std %.3f, smallest top-2 margin %.3e"
      % (out, time.time() - t0, toks[:8].tolist(), float(logits.std()), float((top_val[:, 0] - top_val[:, 1]).min())))
