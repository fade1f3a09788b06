"""Prefill latency of the engine (CUDA-event time from llmc_engine_stats) for a few shapes / prompt lengths.
usage: python profiles/prefill_bench.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as L

for name, shape, T in (("llama-1b", L.SHAPE_LLAMA_1B, 512), ("llama-1b", L.SHAPE_LLAMA_1B, 2048),
                       ("llama-8b", L.SHAPE_LLAMA_8B, 512), ("llama-8b", L.SHAPE_LLAMA_8B, 2048)):
    e = L.Engine(shape=shape, max_slots=1, max_ctx=4096)
    toks = (np.arange(T) * 7919 % 1000 + 32).astype(np.int32)
    ms = []
    for it in range(4):
        s = e.seq_alloc()
        e.prefill(s, toks)
        ms.append(e.stats()["last_prefill_ms"])
        e.seq_free(s)
    print("%s T=%d prefill ms (4 runs, first is cold): %s  -> %.0f tok/s" %
          (name, T, " ".join("%.2f" % m for m in ms), T / min(ms) * 1e3), flush=True)
    del e
