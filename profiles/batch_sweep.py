"""Config 5 (decode batch sweep) on one GPU: tok/s/GPU and fraction of the HBM roofline for batches the
engine serves with one pass over the weights: 1 (persistent kernel), 2..8 (GEMV kernels), 9..128 (wide path:
tcgen05 skinny GEMMs).  usage: python profiles/batch_sweep.py [1b|8b] [max_batch]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc

shape = sys.argv[1] if len(sys.argv) > 1 else "1b"
sid, prompt_len = (llmc.SHAPE_LLAMA_1B, 512) if shape == "1b" else (llmc.SHAPE_LLAMA_8B, 2048)
peak = 6545.0
p = os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")
if os.path.exists(p):
    peak = json.load(open(p))["hbm_gbs"]
steps = 32
max_b = int(sys.argv[2]) if len(sys.argv) > 2 else 128
for kv_flag, kv_name in ((0, "fp32 KV"), (llmc.FLAG_KV_BF16, "bf16 KV")):
    eng = llmc.Engine(shape=sid, seed=0, max_ctx=prompt_len + 512, max_slots=max_b, flags=kv_flag)
    prompt = np.random.default_rng(1).integers(0, 128256, prompt_len)
    slots = []
    for B in [b for b in (1, 2, 4, 8, 16, 32, 64, 128) if b <= max_b]:
        while len(slots) < B:
            s = eng.seq_alloc()
            eng.prefill(s, prompt)
            slots.append(s)
        eng.decode(slots[:B], 2)
        eng.decode(slots[:B], steps)
        st = eng.stats()
        ms = st["last_decode_ms"] / steps
        ctx = eng.seq_len(slots[0]) - steps / 2
        groups = (B + 127) // 128
        algo = st["weight_bytes"] + B * (ctx + 1) * st["kv_bytes_per_token"]          # weights once per step
        moved = groups * st["weight_bytes"] + B * (ctx + 1) * st["kv_bytes_per_token"]  # what this build reads
        print("%s %-8s B=%-3d %.3f ms/step  %8.0f tok/s/GPU  algorithmic %.2f GB -> %.3f of %.0f GB/s (bytes actually "
              "streamed: %.3f)" % (shape, kv_name, B, ms, B / ms * 1e3, algo / 1e9, algo / ms / 1e6 / peak, peak,
                                   moved / ms / 1e6 / peak))
    eng.close()
