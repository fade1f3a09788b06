"""Batch-1 decode of the 1B shape under the ring kernel's variants (grid barriers / tagged-word hand-overs, polling
modes), ms per step at config 2's context (512 -> 768), plus the per-phase trace of block 0.
usage: python profiles/ring_variants.py [trace]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc

trace = len(sys.argv) > 1 and sys.argv[1] == "trace"
variants = [("barriers", {"LLMC_RING_FLOW": "0"}),
            ("barriers, mma.sync slots", {"LLMC_RING_FLOW": "0", "LLMC_RING_MMA": "1"}),
            ("flow, warp sentinel", {"LLMC_RING_FLOW": "1", "LLMC_FLOW_POLL": "0"}),
            ("flow, all threads poll", {"LLMC_RING_FLOW": "1", "LLMC_FLOW_POLL": "1"})]
extra = os.environ.get("RING_VARIANT_TPB")
if extra:
    for t in extra.split(","):
        variants.append(("barriers, %s attention tiles/block" % t, {"LLMC_RING_FLOW": "0", "LLMC_MEGA_TILES_PER_BLOCK": t}))
prompt = np.random.default_rng(1).integers(0, 128256, 512)
ref_toks = None
for name, env in variants:
    for k in ("LLMC_RING_FLOW", "LLMC_FLOW_POLL", "LLMC_RING_POLL", "LLMC_RING_MMA", "LLMC_MEGA_TILES_PER_BLOCK", "LLMC_MEGA_TRACE"):
        os.environ.pop(k, None)
    os.environ.update(env)
    if trace:
        os.environ["LLMC_MEGA_TRACE"] = "1"
    eng = llmc.Engine(shape=llmc.SHAPE_LLAMA_1B, seed=0, max_ctx=1024, max_slots=2)
    best = None
    for rep in range(1 if trace else 3):
        slot = eng.seq_alloc()
        first = eng.prefill(slot, prompt)
        toks, ms, left = [first], 0.0, 255
        while left > 0:
            k = min(64, left) if not trace else (100 if left > 100 else left)   # the trace prints for calls of <= 64 steps
            toks += eng.decode([slot], k)[0].tolist()
            ms += eng.stats()["last_decode_ms"]
            left -= k
        eng.seq_free(slot)
        best = ms / 255 if best is None else min(best, ms / 255)
    if ref_toks is None:
        ref_toks = toks
    same = sum(int(a == b) for a, b in zip(toks, ref_toks))
    print("%-36s %.4f ms/step   tokens equal to the first variant: %d/%d" % (name, best, same, len(toks)), flush=True)
    eng.close()
