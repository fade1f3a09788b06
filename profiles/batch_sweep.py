"""Config 5 (BASELINE.json configs[4]): decode batch sweep 1/8/32/128 (plus 2/4/16/64) on one GPU, for the 1B shape
(context 512 -> 768) and the 8B shape (context 2048 -> 2560): tok/s/GPU and the fraction of the HBM roofline.

Every point starts from FRESH slots (each sequence prefilled with its own prompt of the configured length), decodes
the configured number of steps, and charges each sequence its own context.  Two byte counts are reported:
  * "8(d)":  SURVEY.md section 8(d)'s algorithmic bytes, W + B * (ctx + 1) * kv_tok with kv_tok = 32 KiB (1B) / 128 KiB
             (8B), i.e. a bf16 KV cache -- the figure the roofline fraction is quoted on;
  * "build": the bytes this build streams with the KV cache it was run with (fp32 doubles kv_tok).
usage: python profiles/batch_sweep.py [1b|8b] [max_batch] [kv: fp32|bf16|both] [steps]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc

shape = sys.argv[1] if len(sys.argv) > 1 else "1b"
max_b = int(sys.argv[2]) if len(sys.argv) > 2 else 128
kvs = sys.argv[3] if len(sys.argv) > 3 else "both"
sid, prompt_len, steps, kv_tok = (llmc.SHAPE_LLAMA_1B, 512, 256, 32768.0) if shape == "1b" else (llmc.SHAPE_LLAMA_8B, 2048, 512, 131072.0)
if len(sys.argv) > 4:
    steps = int(sys.argv[4])
peak = 6545.0
p = os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")
if os.path.exists(p):
    peak = json.load(open(p))["hbm_gbs"]
rows = []
for kv_flag, kv_name in ((0, "fp32"), (llmc.FLAG_KV_BF16, "bf16")):
    if kvs not in ("both", kv_name):
        continue
    for B in [b for b in (1, 2, 4, 8, 16, 32, 64, 128) if b <= max_b]:
        eng = llmc.Engine(shape=sid, seed=0, max_ctx=prompt_len + steps + 8, max_slots=B, flags=kv_flag)
        rng = np.random.default_rng(B)
        slots = []
        for _ in range(B):
            s = eng.seq_alloc()
            eng.prefill(s, rng.integers(0, 128256, prompt_len))
            slots.append(s)
        eng.decode(slots, 4)                       # warm-up: graph capture / first cooperative launch
        done, ms = 4, 0.0
        while done < steps:
            k = min(64, steps - done)
            eng.decode(slots, k)
            ms += eng.stats()["last_decode_ms"]
            done += k
        st = eng.stats()
        lens = [eng.seq_len(s) for s in slots]
        assert all(l == prompt_len + steps for l in lens), lens
        timed = steps - 4
        ms_step = ms / timed
        ctx_sum = sum(l - timed / 2.0 for l in lens)          # mean context of each sequence over the timed steps
        algo = st["weight_bytes"] + (ctx_sum + B) * kv_tok
        moved = st["weight_bytes"] + (ctx_sum + B) * st["kv_bytes_per_token"]
        row = {"shape": shape, "kv": kv_name, "B": B, "ms_per_step": ms_step, "tok_s_per_gpu": B / ms_step * 1e3,
               "ctx": "%d -> %d" % (prompt_len, prompt_len + steps), "frac_8d": algo / ms_step / 1e6 / peak,
               "frac_build_bytes": moved / ms_step / 1e6 / peak, "algo_gb": algo / 1e9, "build_gb": moved / 1e9}
        rows.append(row)
        print("%s KV %-4s B=%-3d %.3f ms/step  %8.0f tok/s/GPU  8(d) bytes %.2f GB -> %.3f of %.0f GB/s | this build's bytes "
              "%.2f GB -> %.3f" % (shape, kv_name, B, ms_step, row["tok_s_per_gpu"], algo / 1e9, row["frac_8d"], peak,
                                   moved / 1e9, row["frac_build_bytes"]), flush=True)
        eng.close()
print(json.dumps(rows))
