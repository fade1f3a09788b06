"""Minimal safetensors writer for the checkpoint-loader tests (format: u64 LE header length, JSON header,
raw little-endian tensor bytes) and the Hugging Face Llama tensor names of a model given as oracle tensors."""
import json
import struct

import numpy as np

from oracle import llama_np as lnp
from oracle import ref


def write_safetensors(path, tensors):
    """tensors: {name: (dtype_str, ndarray)}; BF16 arrays are uint16 bit patterns."""
    header, blobs, off = {"__metadata__": {"format": "pt"}}, [], 0
    for name, (dt, arr) in tensors.items():
        raw = np.ascontiguousarray(arr).tobytes()
        header[name] = {"dtype": dt, "shape": list(arr.shape), "data_offsets": [off, off + len(raw)]}
        blobs.append(raw)
        off += len(raw)
    hj = json.dumps(header).encode()
    hj += b" " * (-len(hj) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hj)))
        f.write(hj)
        for b in blobs:
            f.write(b)


def hf_tensors(oracle, dims, dtype="BF16"):
    """The oracle's weights under Hugging Face Llama names (kinds as in oracle/ref.py)."""
    H, I, V = dims["hidden"], dims["inter"], dims["vocab"]
    QD, KD = dims["q_heads"] * dims["head_dim"], dims["kv_heads"] * dims["head_dim"]

    def conv(bits, shape):
        bits = bits.reshape(shape)
        if dtype == "BF16":
            return ("BF16", bits)
        if dtype == "F32":
            return ("F32", lnp.bf16_to_f32(bits).astype(np.float32))
        return ("F16", lnp.bf16_to_f32(bits).astype(np.float16))

    t = {"model.embed_tokens.weight": conv(oracle.tensor(-1, ref.K_EMB), (V, H)),
         "model.norm.weight": conv(oracle.tensor(-1, ref.K_FINAL_NORM), (H,))}
    if not dims["tied"]:
        t["lm_head.weight"] = conv(oracle.tensor(-1, ref.K_LMHEAD), (V, H))
    for l in range(dims["layers"]):
        p = "model.layers.%d." % l
        t[p + "input_layernorm.weight"] = conv(oracle.tensor(l, ref.K_ATTN_NORM), (H,))
        t[p + "self_attn.q_proj.weight"] = conv(oracle.tensor(l, ref.K_WQ), (QD, H))
        t[p + "self_attn.k_proj.weight"] = conv(oracle.tensor(l, ref.K_WK), (KD, H))
        t[p + "self_attn.v_proj.weight"] = conv(oracle.tensor(l, ref.K_WV), (KD, H))
        t[p + "self_attn.o_proj.weight"] = conv(oracle.tensor(l, ref.K_WO), (H, QD))
        t[p + "post_attention_layernorm.weight"] = conv(oracle.tensor(l, ref.K_MLP_NORM), (H,))
        t[p + "mlp.gate_proj.weight"] = conv(oracle.tensor(l, ref.K_WG), (I, H))
        t[p + "mlp.up_proj.weight"] = conv(oracle.tensor(l, ref.K_WU), (I, H))
        t[p + "mlp.down_proj.weight"] = conv(oracle.tensor(l, ref.K_WD), (H, I))
    return t
