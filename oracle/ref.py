"""ctypes binding of the CPU oracle (oracle/llama_ref.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(llm-consensus_b200/) never imports this module.

Parity: unpinned by the reference (it has no local-inference arithmetic; see the
header of llama_ref.cpp).  Pinned instead against oracle/llama_np.py and against
HuggingFace transformers fixtures (tests/test_oracle_pin.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")

# shapes named by BASELINE.json configs
SHAPES = {
    "llama-1b": dict(hidden=2048, layers=16, q_heads=32, kv_heads=8, head_dim=64, inter=8192,
                     vocab=128256, tied=1),
    "llama-8b": dict(hidden=4096, layers=32, q_heads=32, kv_heads=8, head_dim=128, inter=14336,
                     vocab=128256, tied=0),
    # small shapes for fast parity tests (same code paths, every dimension reduced)
    "tiny": dict(hidden=256, layers=2, q_heads=8, kv_heads=2, head_dim=32, inter=512,
                 vocab=1024, tied=1),
    "tiny-untied": dict(hidden=256, layers=2, q_heads=4, kv_heads=4, head_dim=64, inter=768,
                        vocab=1000, tied=0),
    "small-d128": dict(hidden=512, layers=3, q_heads=4, kv_heads=2, head_dim=128, inter=1024,
                       vocab=2048, tied=0),
    # hidden = 2048 shapes with few layers / a small vocabulary: the widths the shared-memory-ring decode
    # kernel handles (rows of 2048 / 4096 / 8192 elements), cheap enough for the CPU oracle
    "wide-1b": dict(hidden=2048, layers=2, q_heads=32, kv_heads=8, head_dim=64, inter=8192,
                    vocab=4096, tied=1),
    "wide-d128": dict(hidden=2048, layers=2, q_heads=32, kv_heads=8, head_dim=128, inter=4096,
                      vocab=2048, tied=0),
    "wide-mha": dict(hidden=2048, layers=1, q_heads=16, kv_heads=16, head_dim=128, inter=2048,
                     vocab=1024, tied=0),
}

K_EMB, K_LMHEAD, K_FINAL_NORM = 0, 1, 2
K_ATTN_NORM, K_WQ, K_WK, K_WV, K_WO, K_MLP_NORM, K_WG, K_WU, K_WD = range(9)


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("hidden", "layers", "q_heads", "kv_heads", "head_dim", "inter", "vocab", "tied")] + \
               [("rope_theta", C.c_float), ("rms_eps", C.c_float), ("init_std", C.c_float),
                ("seed", C.c_uint64), ("max_ctx", C.c_int32),
                # HF rope_type "llama3" (0 = off): factor, low/high frequency factors, original context
                ("rope_scaling", C.c_int32), ("rope_factor", C.c_float), ("rope_low_freq_factor", C.c_float),
                ("rope_high_freq_factor", C.c_float), ("rope_original_ctx", C.c_int32)]


def build(force=False):
    src = os.path.join(_HERE, "llama_ref.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(build())
        l.oref_create.restype = C.c_void_p
        l.oref_create.argtypes = [C.POINTER(Config)]
        l.oref_destroy.argtypes = [C.c_void_p]
        l.oref_reset.argtypes = [C.c_void_p]
        l.oref_len.argtypes = [C.c_void_p]
        l.oref_set_round_mask.argtypes = [C.c_void_p, C.c_int]
        l.oref_truncate.argtypes = [C.c_void_p, C.c_int]
        l.oref_tensor.restype = C.c_int64
        l.oref_tensor.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64]
        l.oref_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        l.oref_generate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        l.oref_argmax.restype = C.c_int32
        l.oref_argmax.argtypes = [C.c_void_p, C.c_int]
        l.oref_num_threads.restype = C.c_int
        _lib = l
    return _lib


def make_config(shape="tiny", seed=0, max_ctx=1024, rope_theta=500000.0, rms_eps=1e-5, init_std=0.02, rope_llama3=None,
                **over):
    """rope_llama3: None, or (factor, low_freq_factor, high_freq_factor, original_max_position_embeddings)."""
    d = dict(SHAPES[shape]) if isinstance(shape, str) else dict(shape)
    d.update(over)
    if rope_llama3 is not None:
        f, lo, hi, oc = rope_llama3
        d.update(rope_scaling=1, rope_factor=f, rope_low_freq_factor=lo, rope_high_freq_factor=hi, rope_original_ctx=int(oc))
    return Config(rope_theta=rope_theta, rms_eps=rms_eps, init_std=init_std, seed=seed, max_ctx=max_ctx, **d)


class Oracle:
    """fp32 CPU forward of the random-init Llama-shaped model."""

    def __init__(self, shape="tiny", seed=0, max_ctx=1024, **kw):
        self.cfg = make_config(shape, seed, max_ctx, **kw)
        self.h = lib().oref_create(C.byref(self.cfg))
        if not self.h:
            raise RuntimeError("oref_create failed")
        self.vocab = self.cfg.vocab

    def close(self):
        if self.h:
            lib().oref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        lib().oref_reset(self.h)

    def set_round_mask(self, mask):
        """Diagnostic: emulate bf16 storage of activations (see llama_ref.cpp); parity tests use mask 0."""
        lib().oref_set_round_mask(self.h, mask)

    def truncate(self, n):
        lib().oref_truncate(self.h, n)

    def __len__(self):
        return lib().oref_len(self.h)

    def tensor(self, layer, kind):
        n = lib().oref_tensor(self.h, layer, kind, None, 0)
        if n < 0:
            raise KeyError((layer, kind))
        out = np.empty(n, np.uint16)
        lib().oref_tensor(self.h, layer, kind, out.ctypes.data, n)
        return out

    def forward(self, tokens, logits="last"):
        t = np.ascontiguousarray(tokens, np.int32)
        mode = {"none": 0, "last": 1, "all": 2}[logits]
        out = None
        if mode == 1:
            out = np.empty(self.vocab, np.float32)
        elif mode == 2:
            out = np.empty((len(t), self.vocab), np.float32)
        rc = lib().oref_forward(self.h, t.ctypes.data, len(t), mode, out.ctypes.data if out is not None else None)
        if rc:
            raise RuntimeError("oref_forward rc=%d" % rc)
        return out

    def generate(self, prompt, n_new, want_logits=False):
        p = np.ascontiguousarray(prompt, np.int32)
        toks = np.empty(n_new, np.int32)
        lg = np.empty((n_new, self.vocab), np.float32) if want_logits else None
        rc = lib().oref_generate(self.h, p.ctypes.data, len(p), n_new, toks.ctypes.data,
                                 lg.ctypes.data if lg is not None else None)
        if rc:
            raise RuntimeError("oref_generate rc=%d" % rc)
        return (toks, lg) if want_logits else toks


def argmax(v):
    v = np.ascontiguousarray(v, np.float32)
    return int(lib().oref_argmax(v.ctypes.data, v.size))


def bf16_bits_to_f32(u16):
    return (u16.astype(np.uint32) << 16).view(np.float32)


def f32_to_bf16_bits(f):
    b = np.ascontiguousarray(f, np.float32).view(np.uint32).astype(np.uint64)
    b = b + 0x7FFF + ((b >> 16) & 1)
    return (b >> 16).astype(np.uint16)


def num_threads():
    return lib().oref_num_threads()
