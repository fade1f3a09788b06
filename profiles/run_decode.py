"""Small driver for ncu captures: 1B engine, 512-token prefill, a few decode steps (graph or not)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
shape = llmc.SHAPE_LLAMA_8B if (len(sys.argv) > 3 and sys.argv[3] == "8b") else llmc.SHAPE_LLAMA_1B
eng = llmc.Engine(shape=shape, seed=0, max_ctx=1024, max_slots=2, flags=flags)
prompt = np.random.default_rng(1).integers(0, 128256, 512)
slot = eng.seq_alloc()
eng.prefill(slot, prompt)
eng.decode([slot], 2)          # warm-up: module load, graph capture, first cooperative launch
eng.decode([slot], steps)
st = eng.stats()
print("prefill_ms", st["last_prefill_ms"], "decode_ms_per_step", st["last_decode_ms"] / steps, "launches", st["kernel_launches"])
