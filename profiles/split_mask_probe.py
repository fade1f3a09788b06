"""Which prefill GEMM operands need the bf16 hi+lo pair for the 2e-2 logit budget?  Config 2 shape (1B, 512-token
prompt), teacher-forced on the oracle's tokens for 48 steps, for every LLMC_PREFILL_SPLIT_MASK (1 QKV, 2 O, 4 gate/up,
8 down).  usage: python profiles/split_mask_probe.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref
import llm_consensus_b200.llmc as L

n_prompt, n_new = 512, 48
o = ref.Oracle("llama-1b", seed=0, max_ctx=1024)
prompt = np.random.default_rng(1).integers(0, 128256, n_prompt)
otoks, olog = o.generate(prompt, n_new, want_logits=True)
for mask in (15, 0, 1, 2, 4, 8, 5, 9, 13, 7, 11, 14):
    os.environ["LLMC_PREFILL_SPLIT_MASK"] = str(mask)
    e = L.Engine(shape=L.SHAPE_LLAMA_1B, seed=0, max_ctx=1024, max_slots=1)
    s = e.seq_alloc()
    nxt, lg = e.prefill(s, prompt, want_logits=True)
    pre_ms = e.stats()["last_prefill_ms"]
    e.seq_free(s); s = e.seq_alloc()
    nxt, lg = e.prefill(s, prompt, want_logits=True)
    pre_ms = e.stats()["last_prefill_ms"]
    errs = [float(np.abs(lg - olog[0]).max())]
    for t in range(1, n_new):
        e.set_next_token(s, otoks[t - 1])
        out, lg = e.decode([s], 1, want_logits=True)
        errs.append(float(np.abs(lg[0] - olog[t]).max()))
    print("mask %2d (QKV %d O %d gate/up %d down %d): prefill logits err %.2e, worst over %d steps %.2e, prefill %.2f ms"
          % (mask, mask & 1, (mask >> 1) & 1, (mask >> 2) & 1, (mask >> 3) & 1, errs[0], n_new, max(errs), pre_ms), flush=True)
    e.close()
