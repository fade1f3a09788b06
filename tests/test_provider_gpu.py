"""The local provider end to end (GPU): the llm-consensus CLI with local model names -- tokenise,
prefill + chunked greedy decode through the C ABI, detokenise, judge prompt, result.json -- against the
same pipeline computed with the CPU oracles (oracle/ref.py for the tokens, oracle/orchestration.py for
the template and the JSON).  This is BASELINE.json configs[2] in miniature: N seeds of one shape, judge
over the concatenated responses."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import orchestration as orc, ref

pytestmark = pytest.mark.gpu
TINY = ref.SHAPES["tiny"]
NEW = 12


@pytest.fixture(scope="module")
def cli(gpu):
    import llm_consensus_b200.build as b
    return b.build_host()[1]


def oracle_answer(seed, prompt_text, n_new=NEW):
    o = ref.Oracle("tiny", seed=seed, max_ctx=8192)
    ids = orc.tokenize(prompt_text, TINY["vocab"])
    toks, lg = o.generate(ids, n_new, want_logits=True)
    margins = [float(np.sort(l)[-1] - np.sort(l)[-2]) for l in lg]
    return orc.detokenize(toks.tolist()), min(margins)


def test_cli_local_fanout_and_judge_matches_oracle_pipeline(cli):
    prompt = "What is 2+2?"
    models = ["local-tiny-s0", "local-tiny-s1", "local-tiny-s2"]
    r = subprocess.run([cli, "--models", ",".join(models), "--judge", "local-tiny-s0", "--max-new-tokens", str(NEW),
                        "--json", prompt], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    doc = json.loads(r.stdout)
    assert doc["prompt"] == prompt and doc["judge"] == "local-tiny-s0"
    assert sorted(x["model"] for x in doc["responses"]) == models
    want, margin = {}, {}
    for i, m in enumerate(models):
        want[m], margin[m] = oracle_answer(i, prompt)
    for x in doc["responses"]:
        assert x["provider"] == "local" and x["latency_ms"] > 0
        if margin[x["model"]] > 1e-3:                          # no numerical tie among the greedy steps
            assert x["content"] == want[x["model"]], x["model"]
    # the judge saw the template rendered over the responses in the order result.json lists them
    rendered = orc.render_judge_prompt(prompt, doc["responses"])
    consensus, jm = oracle_answer(0, rendered)
    if jm > 1e-3:
        assert doc["consensus"] == consensus
    # and the whole document is what Go's encoder would have written for these values
    resps = [dict(x) for x in doc["responses"]]
    assert r.stdout == orc.encode_result(prompt, resps, doc["consensus"], doc["judge"])
    # generated text is printable ASCII that re-tokenises to the same ids
    for x in doc["responses"]:
        ids = orc.tokenize(x["content"], TINY["vocab"])
        assert len(ids) == NEW and orc.detokenize(ids) == x["content"]


def test_cli_single_local_model_passthrough(cli):
    r = subprocess.run([cli, "--models", "local-tiny", "--judge", "local-tiny", "--max-new-tokens", "5", "--seed", "7",
                        "--json", "hello"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    doc = json.loads(r.stdout)
    want, margin = oracle_answer(7, "hello", 5)
    assert doc["consensus"] == doc["responses"][0]["content"]  # one response: judge not called (judge.go:74-79)
    if margin > 1e-3:
        assert doc["consensus"] == want


def test_cli_local_timeout_lands_in_warnings(cli):
    # a 1 s per-model timeout with several seconds' worth of tokens (1B shape, ~1 ms per token): the context
    # is polled between decode chunks, so the query is abandoned and reported like any provider error
    r = subprocess.run([cli, "--models", "local-llama-1b-s0", "--judge", "j", "--stub", "j=openai:0:4",
                        "--max-new-tokens", "7000", "--timeout", "1", "--json", "hi"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 1
    assert r.stderr == "error: running queries: all models failed: [local-llama-1b-s0: context deadline exceeded]\n"


def test_cli_duplicate_model_entries_share_one_engine(cli):
    """runner.go:62 starts one goroutine per --models entry, duplicates included, against the one provider
    initRegistry built for the name (main.go:399-403).  Five entries of one local model (which is also the judge)
    therefore run as five concurrent queries on ONE engine: none may fail for want of a slot, and each answer is
    the model's answer."""
    prompt = "What is 2+2?"
    r = subprocess.run([cli, "--models", ",".join(["local-tiny"] * 5), "--judge", "local-tiny", "--max-new-tokens",
                        str(NEW), "--json", prompt], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    doc = json.loads(r.stdout)
    assert "warnings" not in doc and "failed_models" not in doc
    assert [x["model"] for x in doc["responses"]] == ["local-tiny"] * 5
    want, margin = oracle_answer(0, prompt)
    contents = {x["content"] for x in doc["responses"]}
    if margin > 1e-3:
        assert contents == {want}
    rendered = orc.render_judge_prompt(prompt, doc["responses"])
    consensus, jm = oracle_answer(0, rendered)
    if jm > 1e-3:
        assert doc["consensus"] == consensus


def test_cli_prompt_cache_gives_the_answers_of_a_full_prefill(cli):
    """The provider keeps the last prompt's KV cache per instance: duplicates of a model clone it (no prefill), and
    the judge prompt -- which starts with the template's fixed head -- forks the common prefix.  The whole result.json
    must be what the same run produces with the cache off (LLMC_PROMPT_CACHE=0): clones are bit-identical by
    construction; the judge's forked-prefix prefill may round differently, so its text is only compared when the
    oracle sees no near-tie among its greedy steps."""
    prompt = "Name three prime numbers and say why each of them is prime."
    args = [cli, "--models", "local-tiny,local-tiny,local-tiny-s1,local-tiny", "--judge", "local-tiny", "--max-new-tokens",
            str(NEW), "--json", prompt]
    docs = []
    for cache in ("1", "0"):
        env = dict(os.environ, LLMC_PROMPT_CACHE=cache)
        r = subprocess.run(args, capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0, r.stderr
        docs.append(json.loads(r.stdout))
    on, off = docs
    key = lambda x: (x["model"], x["content"])
    assert sorted(map(key, on["responses"])) == sorted(map(key, off["responses"]))
    same_judge_input = [key(x) for x in on["responses"]] == [key(x) for x in off["responses"]]
    rendered = orc.render_judge_prompt(prompt, on["responses"])
    _, jm = oracle_answer(0, rendered)
    if same_judge_input and jm > 1e-3:
        assert on["consensus"] == off["consensus"]


def test_cli_duplicates_decode_together_on_the_wide_path(cli):
    """Sixteen entries of the 1B model on one GPU: the provider's scheduler advances all sixteen sequences with one
    pass over the weights per step, so the fan-out takes a fraction of sixteen queries run one after another."""
    words = "the quick brown fox jumps over the lazy dog while consensus is computed "
    prompt = (words * 8)[:511] + "?"

    def fanout_ms(n):
        env = dict(os.environ, LLMC_BENCH_REPEAT="3", CUDA_VISIBLE_DEVICES="0")
        r = subprocess.run([cli, "--models", ",".join(["local-llama-1b"] * n), "--judge", "j", "--stub", "j=openai:0:ok",
                            "--max-new-tokens", "256", "--json", prompt], capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stderr[-2000:]
        doc = json.loads(r.stdout)
        assert len(doc["responses"]) == n and "warnings" not in doc
        fans = [float(m.group(1)) for m in re.finditer(r"fan-out ([0-9.]+) ms", r.stderr)]
        return min(fans[1:]), doc

    one, doc1 = fanout_ms(1)
    many, doc16 = fanout_ms(16)
    speedup = 16 * one / many
    print("one query %.1f ms, sixteen concurrent %.1f ms: %.1fx the serial throughput" % (one, many, speedup))
    assert speedup >= 4.0
    same = sum(x["content"] == doc1["responses"][0]["content"] for x in doc16["responses"])
    print("%d of 16 batched answers are character-identical to the single query's" % same)
    assert same >= 12          # batch-1 and batched kernels may part ways at a numerical tie, rarely
