"""Prints the constants held by oracle/go/llamaref/llamaref_test.go (weights of five tensors, eight greedy tokens and
their top logits on the tiny-untied shape, seed 2) from the C++ oracle.  usage: python tests/golden/gen_go_twin_values.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref

o = ref.Oracle("tiny-untied", seed=2, max_ctx=64)
for layer, kind, name in ((-1, ref.K_EMB, "emb"), (-1, ref.K_LMHEAD, "lmhead"), (-1, ref.K_FINAL_NORM, "gf"),
                          (0, ref.K_ATTN_NORM, "g1"), (1, ref.K_WD, "wd")):
    t = o.tensor(layer, kind)
    print(name, len(t), [int(x) for x in t[:4]], int(t[-1]))
prompt = np.array([(37 * i + 11) % 1000 for i in range(12)], dtype=np.int32)
toks, lg = o.generate(prompt, 8, want_logits=True)
print([int(t) for t in toks], [round(float(l.max()), 5) for l in lg],
      "min top-2 margin %.4f" % min(float(np.sort(l)[-1] - np.sort(l)[-2]) for l in lg))
