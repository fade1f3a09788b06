"""Tensor-pipe utilisation of the prefill GEMM (gemm_tcgen05.cu) on the BASELINE shapes.
TFLOP/s counts the algorithmic 2*T*N*K once, also when the A operand is a hi+lo pair (the kernel then issues
twice the MMAs)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_consensus_b200.llmc as llmc
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0}
peak = peaks["bf16_tflops"]
shapes = {"1b": [("qkv", 3072, 2048, 1), ("o", 2048, 2048, 2), ("gateup", 16384, 2048, 3), ("down", 2048, 8192, 2)],
          "8b": [("qkv", 6144, 4096, 1), ("o", 4096, 4096, 2), ("gateup", 28672, 4096, 3), ("down", 4096, 14336, 2)]}
print("model T    op      N      K  split   ms     TFLOP/s  of %.0f (burst)" % peak)
for model, T in (("1b", 512), ("8b", 2048), ("8b", 6400)):
    for name, N, K, epi in shapes[model]:
        for split in (0, 1):
            ms = llmc.op_gemm_bench(T, N, K, epi=epi, split=split, iters=10)
            tf = 2.0 * T * N * K / ms / 1e9
            print("%-5s %-5d %-7s %-6d %-6d %d   %7.3f  %8.1f   %.3f" % (model, T, name, N, K, split, ms, tf, tf / peak))
