"""Per-phase trace (block 0) of the batched ring kernel at the 1B shape, context 512: python profiles/ring_batch_trace.py B
(the cycle counters of the "[ring batch stats]" line are zero unless the library was built with
LLMC_NVCC_EXTRA="-DLLMC_RINGB_STATS" python llm-consensus_b200/build.py --force)"""
import os, sys
import numpy as np
os.environ["LLMC_MEGA_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
eng = llmc.Engine(shape=llmc.SHAPE_LLAMA_1B, seed=0, max_ctx=1024, max_slots=B)
rng = np.random.default_rng(1)
slots = []
for _ in range(B):
    s = eng.seq_alloc()
    eng.prefill(s, rng.integers(0, 128256, 512))
    slots.append(s)
eng.decode(slots, 100)
eng.decode(slots, 60)
print("B=%d: %.4f ms/step" % (B, eng.stats()["last_decode_ms"] / 60))
