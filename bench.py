#!/usr/bin/env python
"""bench.py -- headline benchmark of the local provider's hot path (BASELINE.json configs[1]).

A "step" is one provider query: prefill a 512-token prompt, then greedy-decode 256 tokens, on a
Llama-3.2-1B-shaped random-init bf16 model, one model instance per GPU (the runner's fan-out maps
model i to GPU i with no collective, /root/reference/internal/runner/runner.go:62-95), seeds =
ranks.  Whole-job metric: generated tokens per second over all GPUs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

For N > 1 launch under torchrun (one rank per GPU); torch.distributed is used only for the
barrier and the max-over-ranks reduction -- the measured path is libllmc.so through its C ABI.

--impl reference times the CPU implementation of the same path (the fp32 oracle, all host
threads) on a bounded sample of the workload; see `cpu_baseline.sample`.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PROMPT_LEN, NEW_TOKENS, DECODE_CHUNK = 512, 256, 64
SHAPE_1B = dict(hidden=2048, layers=16, q_heads=32, kv_heads=8, head_dim=64, inter=8192, vocab=128256, tied=1)
METRIC = "decode tok/s (generated tokens / s over all GPUs; the 512-token prefill is inside the timed region)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def cpu_sample(threads_all=True, prompt_len=128, new_tokens=64, steps=3, warmup=0):
    """The fp32 CPU oracle on a bounded sample of the workload (same shape, same prompt/decode ratio)."""
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU legs are meant to use all host threads
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    from oracle import ref
    o = ref.Oracle("llama-1b", seed=0, max_ctx=prompt_len + new_tokens + 8)
    prompt = np.random.default_rng(1).integers(0, 128256, prompt_len)
    times = []
    for i in range(warmup + steps):
        o.reset()
        t0 = time.perf_counter()
        o.generate(prompt, new_tokens)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times)
    return {"value": new_tokens * len(times) / dt, "unit": "tok/s", "cores": ref.num_threads(), "kind": "port",
            "sample": "fp32 C++/OpenMP oracle (oracle/llama_ref.cpp), Llama-3.2-1B shape, %d-token prompt + %d greedy "
                      "tokens per step (the 512+256 workload scaled down, same 2:1 ratio), %d step(s), %.1f s"
                      % (prompt_len, new_tokens, len(times), dt)}, dt / len(times)


OURS_CONFIG = {"workload": "local provider query, Llama-3.2-1B shape random-init bf16, 512-token prompt + 256 greedy "
                           "tokens, batch 1 per GPU (BASELINE.json configs[1]); one model instance (seed = rank) per GPU, "
                           "no collective",
               "shape": "llama-3.2-1b", "prompt": PROMPT_LEN, "new_tokens": NEW_TOKENS, "batch": 1}


def run_reference(args, rank, world):
    """The reference arm: the CPU implementation of the same path (the fp32 oracle port: the reference repository has no
    local inference and Go is not installed) on all host threads.  It runs EXACTLY --steps timed and --warmup untimed
    steps; each step is a bounded sample of the 512 + 256 workload (same shape, same 2:1 prompt/decode ratio), sized
    from the first step's time so that the whole run stays within a few minutes."""
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    budget_s = 240.0
    # probe: one 1/16-scale step tells how fast this host is (its cost is ~1/16 of a quarter-scale step's)
    probe, probe_ms = cpu_sample(prompt_len=32, new_tokens=16, steps=1)
    per_token_s = probe_ms / (32 + 16)
    scale = 4
    for cand in (1, 2, 4, 8, 16):   # the largest sample (smallest divisor) that fits the budget
        if per_token_s * (PROMPT_LEN + NEW_TOKENS) / cand * (steps + warmup) * 1.3 <= budget_s:
            scale = cand
            break
    else:
        scale = 16
    pl, nt = PROMPT_LEN // scale, NEW_TOKENS // scale
    base, ms = cpu_sample(prompt_len=pl, new_tokens=nt, steps=steps, warmup=warmup)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": "tok/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(OURS_CONFIG),
            "sample": "each CPU step is a 1/%d-scale sample of the configured workload: %d-token prompt + %d greedy tokens"
                      % (scale, pl, nt),
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "the reference repository has no local inference (remote HTTP providers only) and Go is not "
                    "installed; the CPU arm is the oracle port of the north-star fp32 forward on all host threads"}
    print(json.dumps(line), flush=True)


class Ranks:
    """The cross-rank plumbing of the N > 1 run: a barrier, a max and a sum over ranks (torch.distributed; NCCL on the
    GPU box, gloo in the CPU test).  The measured path itself has no collective."""

    def __init__(self, world, local_rank=0, backend="nccl"):
        self.world, self.dist, self.device = world, None, "cpu"
        if world > 1:
            import torch
            import torch.distributed as dist
            if backend == "nccl":
                torch.cuda.set_device(local_rank)
                dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
                self.device = "cuda"
            else:
                dist.init_process_group(backend)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def _reduce(self, v, op):
        if self.dist is None:
            return float(v)
        import torch
        t = torch.tensor([v], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return float(t.item())

    def allmax(self, v):
        return self._reduce(v, "MAX")

    def allsum(self, v):
        return self._reduce(v, "SUM")

    def host_signal(self, key):
        """Rank 0 marks `key` in the rendezvous store; host_wait() blocks the other ranks on the HOST until then.  (An
        NCCL barrier would park a spinning kernel on every waiting rank's GPU -- exactly the GPUs the consensus leg of
        rank 0 is about to use.)"""
        if self.dist is not None:
            self.dist.distributed_c10d._get_default_store().set(key, "1")

    def host_wait(self, key):
        if self.dist is not None:
            import datetime
            self.dist.distributed_c10d._get_default_store().wait([key], datetime.timedelta(seconds=1800))

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def aggregate(ranks, dev_pre_ms, dev_dec_ms, wall_s, launches, steps, new_tokens=NEW_TOKENS):
    """Whole-job numbers from this rank's timings: times are the MAX over ranks, work is the SUM over ranks."""
    dev_ms = ranks.allmax(dev_pre_ms + dev_dec_ms)
    out = {"dev_ms": dev_ms, "wall_s": ranks.allmax(wall_s), "dec_ms": ranks.allmax(dev_dec_ms),
           "pre_ms": ranks.allmax(dev_pre_ms), "launches_all": ranks.allsum(float(launches)),
           "tokens_all": new_tokens * steps * ranks.world}
    out["value"] = out["tokens_all"] / (dev_ms / 1e3)
    out["e2e"] = out["tokens_all"] / out["wall_s"]
    return out


def consensus_leg(n_gpus, passes=3):
    """BASELINE.json's first metric, consensus runs/sec, on config 3: 8 seeds of the 1B shape fanned out over the
    first n_gpus GPUs (model i -> GPU i mod n_gpus), judge = model 0's instance on GPU 0 over the rendered template,
    through the CLI mirror of /root/reference/cmd/llm-consensus/main.go:132-163 (fan-out, then judge)."""
    import re
    import tempfile
    cli = os.path.join(ROOT, "llm-consensus_b200", "bin", "llm-consensus")
    if not os.path.exists(cli):
        return {"error": "CLI not built"}
    words = "the quick brown fox jumps over the lazy dog while consensus is computed on eight accelerators "
    prompt = (words * (PROMPT_LEN // len(words) + 1))[:PROMPT_LEN - 1] + "?"
    models = ",".join("local-llama-1b-s%d" % i for i in range(8))
    with tempfile.TemporaryDirectory() as td:
        pf, out = os.path.join(td, "prompt.txt"), os.path.join(td, "result.json")
        open(pf, "w").write(prompt)
        cmd = [cli, "--models", models, "--judge", "local-llama-1b-s0", "--file", pf, "--output", out,
               "--max-new-tokens", str(NEW_TOKENS), "--max-ctx", "8192", "--timeout", "600", "--quiet"]
        # the first n_gpus devices this process can see (an inherited CUDA_VISIBLE_DEVICES is narrowed, not replaced)
        seen = [d for d in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if d.strip()]
        devs = seen[:n_gpus] if len(seen) >= n_gpus else [str(i) for i in range(n_gpus)]
        env = dict(os.environ, LLMC_BENCH_REPEAT=str(passes), CUDA_VISIBLE_DEVICES=",".join(devs))
        try:
            r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        except subprocess.TimeoutExpired:
            return {"error": "CLI timed out"}
        if r.returncode != 0:
            return {"error": "CLI rc=%d: %s" % (r.returncode, (r.stderr or r.stdout)[-300:])}
        runs = [float(m.group(1)) for m in re.finditer(r"run ([0-9.]+) ms", r.stderr)]
        fans = [float(m.group(1)) for m in re.finditer(r"fan-out ([0-9.]+) ms", r.stderr)]
        judges = [float(m.group(1)) for m in re.finditer(r"judge ([0-9.]+) ms", r.stderr)]
        doc = json.load(open(out))
    steady = runs[1:] or runs
    k = len(runs) - len(steady)
    return {"consensus_runs_per_s": len(steady) / (sum(steady) / 1e3), "fanout_ms": sum(fans[k:]) / len(fans[k:]),
            "judge_ms": sum(judges[k:]) / len(judges[k:]), "passes": len(runs), "responses": len(doc["responses"]),
            "config": "BASELINE.json configs[2]: 8 seeds of the Llama-3.2-1B shape over %d GPU(s), 512-char prompt, %d new "
                      "tokens each, judge = model 0 on GPU 0; wall clock inside the CLI, first pass dropped" % (n_gpus, NEW_TOKENS)}


def decode_8b_leg(llmc, device, peak):
    """Batch-1 decode of the 8B shape (BASELINE.json configs[3]'s model), 512-token prompt, 128 steps: ms per step and
    the fraction of the HBM roof on SURVEY section 8(d)'s bytes."""
    try:
        eng = llmc.Engine(shape=llmc.SHAPE_LLAMA_8B, device=device, seed=0, max_ctx=1024, max_slots=1)
    except Exception as ex:   # e.g. not enough free memory next to another tenant
        return {"error": str(ex)[:200]}
    prompt = np.random.default_rng(2).integers(0, 128256, PROMPT_LEN).astype(np.int32)
    slot = eng.seq_alloc()
    eng.prefill(slot, prompt)
    pre = eng.stats()["last_prefill_ms"]
    eng.decode([slot], 8)
    ms, steps = 0.0, 0
    for _ in range(2):
        eng.decode([slot], 64)
        ms += eng.stats()["last_decode_ms"]
        steps += 64
    st = eng.stats()
    eng.close()
    ctx = PROMPT_LEN + 8 + steps / 2.0
    algo = st["weight_bytes"] + (ctx + 1) * 131072.0     # kv_tok = 128 KiB for the 8B shape
    step_ms = ms / steps
    return {"ms_per_step": step_ms, "tok_s": 1e3 / step_ms, "prefill_ms_512": pre,
            "roofline_frac": algo / (step_ms / 1e3) / 1e9 / peak, "algorithmic_bytes_per_step": algo}


def batch_sweep_leg(llmc, device, peak):
    """BASELINE.json configs[4] (decode batch sweep) at the 1B shape on this GPU: B sequences, each prefilled with its
    own 512-token prompt in fresh slots, 64 decode steps (context 512 -> 580); tok/s per GPU and the fraction of the HBM
    roof on SURVEY section 8(d)'s bytes (weights + B x context x 32 KiB).  profiles/batch_sweep.py is the full sweep
    (512 -> 768, both shapes, both KV formats)."""
    out = {}
    for B in (2, 8, 32, 128):
        try:
            eng = llmc.Engine(shape=llmc.SHAPE_LLAMA_1B, device=device, seed=0, max_ctx=640, max_slots=B)
        except Exception as ex:
            out[str(B)] = {"error": str(ex)[:200]}
            continue
        rng = np.random.default_rng(B)
        slots = []
        for _ in range(B):
            sl = eng.seq_alloc()
            eng.prefill(sl, rng.integers(0, 128256, PROMPT_LEN).astype(np.int32))
            slots.append(sl)
        eng.decode(slots, 4)                      # graph capture / first cooperative launch
        ms = 0.0
        for _ in range(4):
            eng.decode(slots, 16)
            ms += eng.stats()["last_decode_ms"]
        st = eng.stats()
        eng.close()
        step_ms = ms / 64
        ctx = PROMPT_LEN + 4 + 32
        algo = st["weight_bytes"] + B * (ctx + 1) * 32768.0
        out[str(B)] = {"ms_per_step": step_ms, "tok_s_per_gpu": B * 1e3 / step_ms,
                       "roofline_frac": algo / (step_ms / 1e3) / 1e9 / peak}
    return out


def run_ours(args, rank, world, local_rank):
    import llm_consensus_b200.llmc as llmc
    ranks = Ranks(world, local_rank, "nccl")
    barrier = ranks.barrier

    if llmc.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- libllmc has no CPU fallback")
    eng = llmc.Engine(shape=llmc.SHAPE_LLAMA_1B, device=local_rank, seed=rank, max_ctx=1024, max_slots=2)
    prompt = np.random.default_rng(1).integers(0, 128256, PROMPT_LEN).astype(np.int32)

    def step():
        """One provider query through the C ABI with host buffers.  Returns (device ms prefill, device ms decode)."""
        slot = eng.seq_alloc()
        eng.prefill(slot, prompt)
        pre = eng.stats()["last_prefill_ms"]
        dec, left = 0.0, NEW_TOKENS - 1
        while left > 0:
            k = min(DECODE_CHUNK, left)
            eng.decode([slot], k)
            dec += eng.stats()["last_decode_ms"]
            left -= k
        eng.seq_free(slot)
        return pre, dec

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = eng.stats()["kernel_launches"]
    barrier()
    t0 = time.perf_counter()
    dev_pre = dev_dec = 0.0
    for _ in range(args.steps):
        p, d = step()
        dev_pre += p
        dev_dec += d
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.summary()
    launches = eng.stats()["kernel_launches"] - launches0
    st = eng.stats()

    agg = aggregate(ranks, dev_pre, dev_dec, wall, launches, args.steps)   # device time: CUDA events on the engine's stream
    dev_ms, wall_s, dec_ms, pre_ms = agg["dev_ms"], agg["wall_s"], agg["dec_ms"], agg["pre_ms"]
    launches_all, tokens_all, value, e2e = agg["launches_all"], agg["tokens_all"], agg["value"], agg["e2e"]
    # roofline of the decode step (one pass of the persistent kernel over all weights): algorithmic bytes / device time
    steps_dec = (NEW_TOKENS - 1) * args.steps
    avg_ctx = PROMPT_LEN + (NEW_TOKENS - 1) / 2.0
    bytes_per_step = st["weight_bytes"] + (avg_ctx + 1) * st["kv_bytes_per_token"]
    step_s = dec_ms / 1e3 / steps_dec
    peak, peak_src = peaks()
    achieved = bytes_per_step / step_s / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "r2_ring_batch_traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp))["dram_bytes_per_step"]
    line = {
        "metric": METRIC, "value": value, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": dict(OURS_CONFIG,
                       l2="no flush needed: each decode step streams 2.47 GB of weights, 20x the 126 MB L2",
                       decode_path="persistent cooperative kernel, weights through a TMA-fed shared-memory ring (batch 1)",
                       kv_cache="fp32", prefill_operands="bf16 hi+lo pairs"),
        "decode_only_tok_s_per_gpu": steps_dec / (dec_ms / 1e3),
        "prefill_ms": pre_ms / args.steps,
        "queries_per_s": args.steps * world / (dev_ms / 1e3),
        "e2e": {"value": e2e, "unit": "tok/s", "h2d_bytes_per_step": PROMPT_LEN * 4 + 8,
                "d2h_bytes_per_step": NEW_TOKENS * 4,
                "what": "wall clock around the llmc_prefill/llmc_decode C-ABI calls with host token buffers "
                        "(H2D of the prompt ids and D2H of the generated ids every step)"},
        "gpu_launches": int(launches_all),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "frac_of_8TBs_nominal": achieved / 8000.0,
                     "kernel": "decode_ring_batch_kernel<64,4,float,1> (the batched ring kernel with one sequence): one persistent "
                               "cooperative kernel runs every decode step of a llmc_decode call (5 phases/layer x 16 + LM "
                               "head per step); figures are per step",
                     "traffic_source": "profiles/r2_ring_batch_traffic.json (ncu --set full of the round-2 build, (dram__bytes_read.sum + "
                                       "dram__bytes_write.sum) / steps)",
                     "algorithmic_bytes_per_launch": bytes_per_step, "avg_launch_us": step_s * 1e6},
        "clocks": clocks,
    }
    eng.close()
    barrier()
    if rank != 0:
        ranks.host_wait("extra_legs_done")      # idle on the host, GPUs free, while rank 0 runs the extra legs
    if rank == 0 and not args.no_extra:
        # extra legs, outside the timed region (other ranks have released their GPUs at the barrier above)
        line["decode_8b"] = decode_8b_leg(llmc, local_rank, peak)
        if world == 1:
            line["batch_sweep_1b"] = batch_sweep_leg(llmc, local_rank, peak)
        line["consensus"] = consensus_leg(world)
        for k in ("consensus_runs_per_s", "fanout_ms", "judge_ms"):
            line[k] = line["consensus"].get(k)
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"], _ = cpu_sample()
    if rank == 0:
        print(json.dumps(line), flush=True)
        ranks.host_signal("extra_legs_done")
    barrier()
    ranks.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the 8B decode, batch sweep (config 5) and consensus (config 3) legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
