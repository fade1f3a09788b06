"""One prefill of T tokens (for ncu launch lists).  usage: python profiles/run_prefill.py [1b|8b] [T]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as L

shape = L.SHAPE_LLAMA_8B if (len(sys.argv) > 1 and sys.argv[1] == "8b") else L.SHAPE_LLAMA_1B
T = int(sys.argv[2]) if len(sys.argv) > 2 else 512
e = L.Engine(shape=shape, max_slots=1, max_ctx=T + 64)
toks = (np.arange(T) * 7919 % 1000 + 32).astype(np.int32)
s = e.seq_alloc()
e.prefill(s, toks)
print("prefill ms", e.stats()["last_prefill_ms"])
