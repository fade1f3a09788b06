"""Configs 3 / 4 end to end through the CLI mirror (llm-consensus_b200/bin/llm-consensus): N local models with
seeds 0..N-1 (model i -> GPU i mod #GPUs), judge = model 0's instance on GPU 0, LLMC_BENCH_REPEAT passes.
usage: python profiles/consensus_bench.py [1b|8b] [n_models] [passes]
Prints the CLI's per-pass timings and consensus runs/sec (steady state = passes after the first)."""
import os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
shape = sys.argv[1] if len(sys.argv) > 1 else "1b"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 4
prompt_len, new_tokens, max_ctx = (512, 256, 32768) if shape == "1b" else (2048, 512, 49152)
words = "the quick brown fox jumps over the lazy dog while consensus is computed on eight accelerators "
prompt = (words * (prompt_len // len(words) + 1))[:prompt_len - 1] + "?"
models = ",".join("local-llama-%s-s%d" % (shape, i) for i in range(n))
with tempfile.TemporaryDirectory() as td:
    pf = os.path.join(td, "prompt.txt")
    open(pf, "w").write(prompt)
    out = os.path.join(td, "result.json")
    cmd = [os.path.join(ROOT, "llm-consensus_b200", "bin", "llm-consensus"), "--models", models, "--judge",
           "local-llama-%s-s0" % shape, "--file", pf, "--output", out, "--max-new-tokens", str(new_tokens), "--max-ctx",
           str(max_ctx), "--timeout", "900", "--quiet"]
    env = dict(os.environ, LLMC_BENCH_REPEAT=str(passes))
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=3000)
    sys.stderr.write(r.stderr)
    if r.returncode != 0:
        sys.exit("CLI failed rc=%d: %s" % (r.returncode, r.stdout[-500:]))
    runs = [float(m.group(1)) for m in re.finditer(r"run ([0-9.]+) ms", r.stderr)]
    fans = [float(m.group(1)) for m in re.finditer(r"fan-out ([0-9.]+) ms", r.stderr)]
    judges = [float(m.group(1)) for m in re.finditer(r"judge ([0-9.]+) ms", r.stderr)]
    steady = runs[1:] or runs
    import json
    doc = json.load(open(out))
    print(json.dumps({"config": "consensus %s x%d models, prompt %d tokens, %d new tokens, judge on GPU 0" %
                      (shape, n, prompt_len, new_tokens), "passes": len(runs),
                      "run_ms": runs, "fanout_ms": fans, "judge_ms": judges,
                      "runs_per_sec_steady": len(steady) / (sum(steady) / 1e3),
                      "fanout_tok_s_all_gpus": n * new_tokens / (min(fans) / 1e3),
                      "responses": len(doc["responses"]), "judge_prompt_chars": None,
                      "consensus_chars": len(doc["consensus"])}))
