"""Generate the golden fixtures that pin the CPU oracle (run once, in the build container):

    python tests/golden/gen_hf_fixtures.py

* hf_<shape>.npz : logits of HuggingFace transformers' LlamaForCausalLM (CPU, fp32) loaded with
  the weights of the oracle's counter RNG, on a fixed token sequence.  transformers is the
  independent implementation of the Llama-3 architecture that the oracle restates; the
  reference repository itself contains no arithmetic to take vectors from.
* rng_kat.json   : known-answer bf16 bit patterns of the weight RNG, produced by the numpy
  restatement (oracle/llama_np.py).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import ref, llama_np as lnp  # noqa: E402


LLAMA3_ROPE = (8.0, 1.0, 4.0, 64)   # factor, low_freq_factor, high_freq_factor, original_max_position_embeddings


def hf_logits(name, seed, n_tok, rope_llama3=None):
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    s = ref.SHAPES[name]
    cfg = LlamaConfig(vocab_size=s["vocab"], hidden_size=s["hidden"], intermediate_size=s["inter"],
                      num_hidden_layers=s["layers"], num_attention_heads=s["q_heads"],
                      num_key_value_heads=s["kv_heads"], head_dim=s["head_dim"], rms_norm_eps=1e-5,
                      rope_theta=500000.0, tie_word_embeddings=bool(s["tied"]), attention_bias=False,
                      mlp_bias=False, max_position_embeddings=256)
    cfg.rope_parameters = {"rope_type": "default", "rope_theta": 500000.0}
    if rope_llama3 is not None:   # the scaling real Llama-3.1 / 3.2 checkpoints declare in config.json
        f, lo, hi, oc = rope_llama3
        cfg.rope_parameters = {"rope_type": "llama3", "rope_theta": 500000.0, "factor": f, "low_freq_factor": lo,
                               "high_freq_factor": hi, "original_max_position_embeddings": oc}
    hf = LlamaForCausalLM(cfg).float().eval()
    H, D = s["hidden"], s["head_dim"]
    QD, KD, I, V = s["q_heads"] * D, s["kv_heads"] * D, s["inter"], s["vocab"]

    def T(layer, kind, rows, cols):
        tid = lnp.tensor_id(layer, kind)
        a = lnp.gain(seed, tid, cols) if rows == 0 else lnp.matrix(seed, tid, rows, cols)
        return torch.from_numpy(np.ascontiguousarray(a))

    sd = {"model.embed_tokens.weight": T(-1, 0, V, H), "model.norm.weight": T(-1, 2, 0, H),
          "lm_head.weight": T(-1, 0 if s["tied"] else 1, V, H)}
    for l in range(s["layers"]):
        p = "model.layers.%d." % l
        sd.update({p + "input_layernorm.weight": T(l, 0, 0, H), p + "self_attn.q_proj.weight": T(l, 1, QD, H),
                   p + "self_attn.k_proj.weight": T(l, 2, KD, H), p + "self_attn.v_proj.weight": T(l, 3, KD, H),
                   p + "self_attn.o_proj.weight": T(l, 4, H, QD), p + "post_attention_layernorm.weight": T(l, 5, 0, H),
                   p + "mlp.gate_proj.weight": T(l, 6, I, H), p + "mlp.up_proj.weight": T(l, 7, I, H),
                   p + "mlp.down_proj.weight": T(l, 8, H, I)})
    res = hf.load_state_dict(sd, strict=False)
    assert not res.missing_keys and not res.unexpected_keys, res
    toks = np.random.default_rng(1234).integers(0, V, n_tok).astype(np.int32)
    with torch.no_grad():
        lg = hf(torch.from_numpy(toks.astype(np.int64))[None]).logits[0].numpy().astype(np.float32)
    return toks, lg


def main():
    for name, seed, n_tok in [("tiny", 5, 40), ("small-d128", 7, 24), ("tiny-untied", 9, 24)]:
        toks, lg = hf_logits(name, seed, n_tok)
        np.savez_compressed(os.path.join(HERE, "hf_%s.npz" % name), tokens=toks, logits=lg, seed=seed)
        print(name, lg.shape, float(lg.std()))
    # RoPE frequency scaling (rope_type "llama3"): 200 positions against an original context of 64, so the three
    # bands (kept / interpolated / stretched) all occur among the 16 frequencies of the tiny shape
    toks, lg = hf_logits("tiny", 6, 200, LLAMA3_ROPE)
    np.savez_compressed(os.path.join(HERE, "hf_tiny_rope_llama3.npz"), tokens=toks, logits=lg, seed=6,
                        rope_llama3=np.asarray(LLAMA3_ROPE, np.float64))
    print("tiny + llama3 rope", lg.shape, float(lg.std()))
    kat = {}
    for seed, layer, kind, rows, cols in [(0, -1, 0, 4, 8), (0, 3, 1, 2, 16), (12345, 0, 8, 3, 8), (7, 15, 6, 2, 8)]:
        m = lnp.matrix(seed, lnp.tensor_id(layer, kind), rows, cols)
        kat["m:%d:%d:%d:%d:%d" % (seed, layer, kind, rows, cols)] = lnp.bf16_rne_bits(m).ravel().tolist()
    for seed, layer, kind, n in [(0, -1, 2, 8), (99, 2, 5, 8)]:
        g = lnp.gain(seed, lnp.tensor_id(layer, kind), n)
        kat["g:%d:%d:%d:%d" % (seed, layer, kind, n)] = lnp.bf16_rne_bits(g).tolist()
    json.dump(kat, open(os.path.join(HERE, "rng_kat.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
