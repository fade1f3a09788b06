"""Checkpoint loading (llmc_config.weights_path / --weights): a safetensors file with Hugging Face Llama tensor
names, written from the ORACLE's weights for seed S, loaded into an engine created with a DIFFERENT seed, must
give the oracle's tensors bit for bit and the oracle's logits within the parity tolerance."""
import os
import subprocess

import numpy as np
import pytest

import sys

from oracle import ref

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from _safetensors import hf_tensors, write_safetensors  # noqa: E402

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DIMS = ref.SHAPES


def check_engine(gpu, o, dims, path):
    e = gpu.Engine(dims=dims, seed=987654, max_ctx=128, max_slots=1, weights_path=path)
    for layer in [-1] + list(range(dims["layers"])):
        kinds = ([0, 2] if dims["tied"] else [0, 1, 2]) if layer < 0 else list(range(9))
        for kind in kinds:
            assert np.array_equal(e.debug_tensor(layer, kind), o.tensor(layer, kind)), (layer, kind)
    toks = np.random.default_rng(3).integers(0, dims["vocab"], 40)
    want = o.forward(toks, "last")
    slot = e.seq_alloc()
    _, lg = e.prefill(slot, toks, want_logits=True)
    assert np.abs(lg - want).max() <= 2e-2
    e.close()


@pytest.mark.parametrize("name,dtype", [("tiny", "BF16"), ("tiny-untied", "BF16"), ("small-d128", "F32")])
def test_load_single_file(gpu, tmp_path, name, dtype):
    dims = DIMS[name]
    o = ref.Oracle(name, seed=31, max_ctx=128)
    t = hf_tensors(o, dims, dtype)
    path = str(tmp_path / "model.safetensors")
    write_safetensors(path, t)
    check_engine(gpu, o, dims, path)


def test_load_sharded_directory(gpu, tmp_path):
    name = "tiny-untied"
    dims = DIMS[name]
    o = ref.Oracle(name, seed=32, max_ctx=128)
    t = hf_tensors(o, dims)
    names = sorted(t)
    d = tmp_path / "ckpt"
    d.mkdir()
    write_safetensors(str(d / "model-00001-of-00002.safetensors"), {n: t[n] for n in names[::2]})
    write_safetensors(str(d / "model-00002-of-00002.safetensors"), {n: t[n] for n in names[1::2]})
    (d / "config.json").write_text("{}")     # other files in the directory are ignored
    check_engine(gpu, o, dims, str(d))


def test_f16_is_rounded_to_bf16(gpu, tmp_path):
    dims = DIMS["tiny"]
    rng = np.random.default_rng(0)
    o = ref.Oracle("tiny", seed=33, max_ctx=64)
    t = hf_tensors(o, dims)
    f16 = (rng.standard_normal((dims["vocab"], dims["hidden"])) * 0.02).astype(np.float16)
    t["model.embed_tokens.weight"] = ("F16", f16)
    path = str(tmp_path / "m.safetensors")
    write_safetensors(path, t)
    e = gpu.Engine(dims=dims, seed=1, max_ctx=64, max_slots=1, weights_path=path)
    from oracle import llama_np as lnp
    assert np.array_equal(e.debug_tensor(-1, 0), lnp.bf16_rne_bits(f16.astype(np.float32)).ravel())
    e.close()


def test_errors_name_the_problem(gpu, tmp_path):
    dims = DIMS["tiny"]
    o = ref.Oracle("tiny", seed=34, max_ctx=64)
    t = hf_tensors(o, dims)
    missing = dict(t)
    del missing["model.layers.1.mlp.up_proj.weight"]
    p1 = str(tmp_path / "missing.safetensors")
    write_safetensors(p1, missing)
    with pytest.raises(gpu.LlmcError) as ei:
        gpu.Engine(dims=dims, seed=1, max_ctx=64, max_slots=1, weights_path=p1)
    assert "model.layers.1.mlp.up_proj.weight" in str(ei.value)
    bad = dict(t)
    bad["model.norm.weight"] = ("BF16", np.zeros(dims["hidden"] + 8, np.uint16))
    p2 = str(tmp_path / "shape.safetensors")
    write_safetensors(p2, bad)
    with pytest.raises(gpu.LlmcError) as ei:
        gpu.Engine(dims=dims, seed=1, max_ctx=64, max_slots=1, weights_path=p2)
    assert "model.norm.weight" in str(ei.value) and "shape" in str(ei.value)
    with pytest.raises(gpu.LlmcError) as ei:
        gpu.Engine(dims=dims, seed=1, max_ctx=64, max_slots=1, weights_path=str(tmp_path / "nope.safetensors"))
    assert "nope.safetensors" in str(ei.value)
    p3 = str(tmp_path / "garbage.safetensors")
    open(p3, "wb").write(b"\xff" * 64)
    with pytest.raises(gpu.LlmcError):
        gpu.Engine(dims=dims, seed=1, max_ctx=64, max_slots=1, weights_path=p3)


def test_cli_weights_flag(gpu, tmp_path):
    """--weights through the CLI mirror: local-tiny loaded from a checkpoint written for seed 35 answers exactly
    like the seeded model local-tiny-s35 (same weights by construction)."""
    dims = DIMS["tiny"]
    o = ref.Oracle("tiny", seed=35, max_ctx=64)
    path = str(tmp_path / "tiny.safetensors")
    write_safetensors(path, hf_tensors(o, dims))
    exe = os.path.join(ROOT, "llm-consensus_b200", "bin", "llm-consensus")
    outs = []
    for extra, model in ((["--weights", path], "local-tiny"), ([], "local-tiny-s35")):
        out = str(tmp_path / ("r_%d.json" % len(outs)))
        r = subprocess.run([exe, "--models", model, "--judge", model, "--output", out, "--max-new-tokens", "12", "--quiet"]
                           + extra + ["hello checkpoint"], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        import json
        outs.append(json.load(open(out))["responses"][0]["content"])
    assert outs[0] == outs[1] and len(outs[0]) > 0


ROPE3 = (8.0, 1.0, 4.0, 64)       # factor, low_freq_factor, high_freq_factor, original_max_position_embeddings


def test_rope_scaling_llama3_engine_matches_oracle(gpu):
    """llmc_config.rope_scaling = LLMC_ROPE_LLAMA3: prefill logits at 120 positions and a few decode steps against
    the oracle with the same scaling (itself pinned to transformers by tests/golden/hf_tiny_rope_llama3.npz)."""
    dims = DIMS["tiny"]
    o = ref.Oracle("tiny", seed=41, max_ctx=256, rope_llama3=ROPE3)
    e = gpu.Engine(dims=dims, seed=41, max_ctx=256, max_slots=1, rope_llama3=ROPE3)
    toks = np.random.default_rng(8).integers(0, dims["vocab"], 120)
    otoks, olog = o.generate(toks, 6, want_logits=True)
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, toks, want_logits=True)
    assert np.abs(lg - olog[0]).max() <= 2e-2
    plain = ref.Oracle("tiny", seed=41, max_ctx=256).forward(toks, "last")
    assert np.abs(lg - plain).max() > 2e-2         # the scaling is visible at this length
    for s_ in range(1, 6):
        e.set_next_token(slot, int(otoks[s_ - 1]))
        _, lgd = e.decode([slot], 1, want_logits=True)
        assert np.abs(lgd[0] - olog[s_]).max() <= 2e-2
    e.close()


def test_config_json_beside_the_checkpoint_sets_rope_scaling(gpu, tmp_path):
    """A checkpoint directory whose config.json declares rope_scaling of type llama3 (as Llama-3.1 / 3.2 do) is run
    with that scaling; a type the library does not implement is refused instead of loading and computing something
    else."""
    import json
    dims = DIMS["tiny"]
    o = ref.Oracle("tiny", seed=42, max_ctx=256, rope_llama3=ROPE3)
    d = tmp_path / "ckpt"
    d.mkdir()
    write_safetensors(str(d / "model.safetensors"), hf_tensors(o, dims))
    cfg = {"architectures": ["LlamaForCausalLM"], "rope_theta": 500000.0, "nested": {"rope_scaling": "decoy"},
           "rope_scaling": {"factor": ROPE3[0], "low_freq_factor": ROPE3[1], "high_freq_factor": ROPE3[2],
                            "original_max_position_embeddings": ROPE3[3], "rope_type": "llama3"}}
    (d / "config.json").write_text(json.dumps(cfg, indent=2))
    toks = np.random.default_rng(9).integers(0, dims["vocab"], 120)
    want = o.forward(toks, "last")
    for path in (str(d), str(d / "model.safetensors")):      # directory, or the file inside it
        e = gpu.Engine(dims=dims, seed=1, max_ctx=256, max_slots=1, weights_path=path)
        slot = e.seq_alloc()
        _, lg = e.prefill(slot, toks, want_logits=True)
        assert np.abs(lg - want).max() <= 2e-2
        e.close()
    cfg["rope_scaling"] = {"rope_type": "yarn", "factor": 4.0}
    (d / "config.json").write_text(json.dumps(cfg))
    with pytest.raises(gpu.LlmcError) as ei:
        gpu.Engine(dims=dims, seed=1, max_ctx=256, max_slots=1, weights_path=str(d))
    assert "yarn" in str(ei.value) and "rope_scaling" in str(ei.value)
    cfg["rope_scaling"] = None
    (d / "config.json").write_text(json.dumps(cfg))
    e = gpu.Engine(dims=dims, seed=1, max_ctx=256, max_slots=1, weights_path=str(d))      # null: plain RoPE
    slot = e.seq_alloc()
    _, lg = e.prefill(slot, toks, want_logits=True)
    o2 = ref.Oracle("tiny", seed=42, max_ctx=256)
    assert np.abs(lg - o2.forward(toks, "last")).max() <= 2e-2
    e.close()
