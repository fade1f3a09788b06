import sys, numpy as np
sys.path.insert(0,'/root/repo')
import llm_consensus_b200.llmc as L
e = L.Engine(shape=L.SHAPE_LLAMA_1B, seed=0, max_ctx=4608, max_slots=1)
p = np.random.default_rng(1).integers(0, 128256, 4100)
s = e.seq_alloc(); e.prefill(s, p); e.decode([s], 2); e.decode([s], 128)
print("ctx 4100 decode ms/step", e.stats()["last_decode_ms"]/128)
