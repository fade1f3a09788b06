"""Batched decode driver for ncu launch lists: python profiles/run_decode_batch.py B [steps] [1b|8b]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
shape = llmc.SHAPE_LLAMA_8B if (len(sys.argv) > 3 and sys.argv[3] == "8b") else llmc.SHAPE_LLAMA_1B
eng = llmc.Engine(shape=shape, seed=0, max_ctx=1024, max_slots=B)
prompt = np.random.default_rng(1).integers(0, 128256, 512)
slots = []
for _ in range(B):
    s = eng.seq_alloc()
    eng.prefill(s, prompt)
    slots.append(s)
eng.decode(slots, 2)
eng.decode(slots, steps)
st = eng.stats()
print("B", B, "decode_ms_per_step", st["last_decode_ms"] / steps)
