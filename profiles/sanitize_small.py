"""Small driver for compute-sanitizer (memcheck / racecheck / synccheck): every kernel family of the decode and
prefill paths on small shapes, a few steps each.
  wide-1b : prefill (gemm_tc_kernel, prefill attention, rmsnorm / rope kernels) + batch-1 decode (decode_ring_kernel)
  tiny    : batch-1 decode (decode_mega_kernel), batch 4 (gemm_decode_kernel, attn_decode_kernel, decode_wide kernels),
            batch 40 (gemm_tc_kernel in decode mode with split-K)
usage: compute-sanitizer --tool memcheck python profiles/sanitize_small.py [ring|mega|wide4|wide40|all]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc
from oracle import ref   # shape table only

what = sys.argv[1] if len(sys.argv) > 1 else "all"


def run(name, batch, n_prompt, steps, flags=0):
    dims = ref.SHAPES[name]
    e = llmc.Engine(dims=dims, seed=3, max_ctx=128, max_slots=max(2, batch), flags=flags)
    rng = np.random.default_rng(batch)
    slots = []
    for _ in range(batch):
        s = e.seq_alloc()
        e.prefill(s, rng.integers(0, dims["vocab"], n_prompt))
        slots.append(s)
    toks = e.decode(slots, steps)
    print("%s batch %d: %d steps ok, first tokens %s" % (name, batch, steps, toks[:, 0].tolist()[:4]), flush=True)
    e.close()


if what in ("ring", "all"):
    run("wide-1b", 1, 40, 3)
if what in ("mega", "all"):
    run("tiny", 1, 40, 3)
if what in ("wide4", "all"):
    run("tiny", 4, 20, 2)
if what in ("wide40", "all"):
    run("tiny", 40, 12, 2)
