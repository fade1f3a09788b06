"""ctypes binding of libllmc.so (include/llmc.h, include/llmc_ops.h).

Thin by design: the product is the CUDA library; this file only marshals numpy buffers
across the C ABI for the Python-side callers (tests, bench.py, the provider mirror).
There is no fallback: if the library is missing, loading raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LLMC_LIB_PATH") or os.path.join(_HERE, "lib", "libllmc.so")   # (override: A/B timing of two builds)

SHAPE_CUSTOM, SHAPE_LLAMA_1B, SHAPE_LLAMA_8B = 0, 1, 2
FLAG_NO_GRAPH, FLAG_NO_PDL, FLAG_KV_BF16, FLAG_PREFILL_BF16 = 1, 2, 4, 8
FLAG_NO_MEGAKERNEL, FLAG_NO_RING, FLAG_RING_FLOW, FLAG_NO_RING_BATCH, FLAG_RING_MMA = 16, 32, 64, 128, 256
OK, ERR_INVALID, ERR_CUDA, ERR_OOM, ERR_STATE, ERR_LIMIT = 0, -1, -2, -3, -4, -5

# every symbol the headers declare (checked by tests/test_abi_symbols.py)
ENGINE_SYMBOLS = [
    "llmc_device_count", "llmc_engine_create", "llmc_engine_destroy", "llmc_seq_alloc", "llmc_seq_free",
    "llmc_seq_len", "llmc_seq_fork", "llmc_prefill", "llmc_decode", "llmc_seq_set_next_token", "llmc_seq_get_next_token",
    "llmc_engine_stats", "llmc_last_error", "llmc_version",
]
OPS_SYMBOLS = [
    "llmc_ops_last_error", "llmc_op_gemv", "llmc_op_gemv_swiglu", "llmc_op_lmhead", "llmc_op_attn_decode",
    "llmc_op_gemm", "llmc_op_gemm_bench", "llmc_op_rmsnorm_rows", "llmc_op_prefill_attention", "llmc_debug_tensor", "llmc_debug_kv",
    "llmc_lmhead_argmax_noalloc", "llmc_noalloc_last_error",
]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("shape", C.c_int32)] + \
               [(n, C.c_int32) for n in ("hidden", "layers", "q_heads", "kv_heads", "head_dim", "inter", "vocab", "tied")] + \
               [("rope_theta", C.c_float), ("rms_eps", C.c_float), ("init_std", C.c_float), ("seed", C.c_uint64),
                ("weights_path", C.c_char_p)] + \
               [(n, C.c_int32) for n in ("max_slots", "max_ctx", "page_size", "prefill_chunk", "prefill_impl", "flags")] + \
               [("rope_scaling", C.c_int32), ("rope_factor", C.c_float), ("rope_low_freq_factor", C.c_float),
                ("rope_high_freq_factor", C.c_float), ("rope_original_ctx", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("decode_steps", C.c_uint64), ("prefill_tokens", C.c_uint64),
                ("weight_bytes", C.c_double), ("kv_bytes_per_token", C.c_double), ("last_prefill_ms", C.c_float),
                ("last_decode_ms", C.c_float), ("graph_enabled", C.c_int32), ("pdl_enabled", C.c_int32),
                ("sm_count", C.c_int32), ("cc_major", C.c_int32), ("cc_minor", C.c_int32)]


class LlmcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("llmc error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError("libllmc.so is not built (%s); run `python llm-consensus_b200/build.py` -- there is no "
                          "fallback implementation" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        vp, i32, f32p = C.c_void_p, C.c_int32, C.c_void_p
        l.llmc_device_count.restype = C.c_int
        l.llmc_engine_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
        l.llmc_engine_destroy.argtypes = [vp]
        l.llmc_engine_destroy.restype = None
        l.llmc_seq_alloc.argtypes = [vp, C.POINTER(i32)]
        l.llmc_seq_free.argtypes = [vp, i32]
        l.llmc_seq_len.argtypes = [vp, i32, C.POINTER(i32)]
        l.llmc_seq_fork.argtypes = [vp, i32, i32, C.POINTER(i32)]
        l.llmc_prefill.argtypes = [vp, i32, vp, i32, C.POINTER(i32), f32p]
        l.llmc_decode.argtypes = [vp, vp, i32, i32, vp, f32p]
        l.llmc_seq_set_next_token.argtypes = [vp, i32, i32]
        l.llmc_seq_get_next_token.argtypes = [vp, i32, C.POINTER(i32)]
        l.llmc_engine_stats.argtypes = [vp, C.POINTER(Stats)]
        l.llmc_last_error.argtypes = [vp]
        l.llmc_last_error.restype = C.c_char_p
        l.llmc_version.restype = C.c_char_p
        l.llmc_ops_last_error.restype = C.c_char_p
        l.llmc_op_gemv.argtypes = [C.c_int] * 4 + [vp, vp, C.c_int, vp, C.c_float, C.c_int, vp]
        l.llmc_op_gemv_swiglu.argtypes = [C.c_int] * 4 + [vp, vp, vp, vp, C.c_float, vp]
        l.llmc_op_lmhead.argtypes = [C.c_int] * 4 + [vp, vp, vp, C.c_float, vp, vp]
        l.llmc_op_attn_decode.argtypes = [C.c_int] * 9 + [vp] * 5
        l.llmc_op_gemm.argtypes = [C.c_int] * 6 + [vp] * 5
        l.llmc_op_gemm_bench.argtypes = [C.c_int] * 8 + [vp]
        l.llmc_op_rmsnorm_rows.argtypes = [C.c_int] * 3 + [vp, vp, C.c_float, vp, vp]
        l.llmc_op_prefill_attention.argtypes = [C.c_int] * 7 + [C.c_float, C.c_int] + [vp] * 8 + [C.c_int]
        l.llmc_debug_tensor.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int64]
        l.llmc_debug_tensor.restype = C.c_int64
        l.llmc_debug_kv.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp]
        l.llmc_lmhead_argmax_noalloc.argtypes = [C.c_int] * 4 + [vp, vp, vp, C.c_float, vp, vp, vp]
        l.llmc_noalloc_last_error.restype = C.c_char_p
        _lib = l
    return _lib


def device_count():
    return lib().llmc_device_count()


def _ptr(a):
    return a.ctypes.data if a is not None else None


class Engine:
    """One model instance on one GPU (llmc_engine)."""

    def __init__(self, shape=SHAPE_LLAMA_1B, device=0, seed=0, max_slots=8, max_ctx=8192, page_size=64,
                 prefill_chunk=2048, prefill_impl=0, flags=0, dims=None, rope_theta=0.0, rms_eps=0.0, init_std=0.0,
                 weights_path=None, rope_llama3=None):
        """rope_llama3: None, or (factor, low_freq_factor, high_freq_factor, original_max_position_embeddings)."""
        cfg = Config(device=device, shape=shape, seed=seed, max_slots=max_slots, max_ctx=max_ctx,
                     page_size=page_size, prefill_chunk=prefill_chunk, prefill_impl=prefill_impl, flags=flags,
                     rope_theta=rope_theta, rms_eps=rms_eps, init_std=init_std,
                     weights_path=weights_path.encode() if weights_path else None)
        if rope_llama3 is not None:
            cfg.rope_scaling, cfg.rope_factor, cfg.rope_low_freq_factor, cfg.rope_high_freq_factor = 1, *map(float, rope_llama3[:3])
            cfg.rope_original_ctx = int(rope_llama3[3])
        if dims:
            cfg.shape = SHAPE_CUSTOM
            for k, v in dims.items():
                setattr(cfg, k, v)
        self.h = C.c_void_p()
        rc = lib().llmc_engine_create(C.byref(cfg), C.byref(self.h))
        if rc != OK:
            self.h = None
            raise LlmcError(rc, lib().llmc_last_error(None).decode())
        if dims:
            self.vocab = dims["vocab"]
        else:
            self.vocab = 128256

    def _check(self, rc):
        if rc != OK:
            raise LlmcError(rc, lib().llmc_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            lib().llmc_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def seq_alloc(self):
        s = C.c_int32(-1)
        self._check(lib().llmc_seq_alloc(self.h, C.byref(s)))
        return s.value

    def seq_free(self, slot):
        self._check(lib().llmc_seq_free(self.h, slot))

    def seq_len(self, slot):
        n = C.c_int32(0)
        self._check(lib().llmc_seq_len(self.h, slot, C.byref(n)))
        return n.value

    def prefill(self, slot, tokens, want_logits=False):
        t = np.ascontiguousarray(tokens, np.int32)
        nxt = C.c_int32(-1)
        lg = np.empty(self.vocab, np.float32) if want_logits else None
        self._check(lib().llmc_prefill(self.h, slot, t.ctypes.data, len(t), C.byref(nxt), _ptr(lg)))
        return (nxt.value, lg) if want_logits else nxt.value

    def seq_fork(self, src, n_tokens=None):
        """New slot holding the first n_tokens (default: all) cached tokens of slot src (include/llmc.h: llmc_seq_fork)."""
        s = C.c_int32(-1)
        n = self.seq_len(src) if n_tokens is None else int(n_tokens)
        self._check(lib().llmc_seq_fork(self.h, int(src), n, C.byref(s)))
        return s.value

    def decode(self, slots, nsteps, want_logits=False):
        s = np.ascontiguousarray(slots, np.int32)
        out = np.empty((len(s), nsteps), np.int32)
        lg = np.empty((len(s), self.vocab), np.float32) if want_logits else None
        self._check(lib().llmc_decode(self.h, s.ctypes.data, len(s), nsteps, out.ctypes.data, _ptr(lg)))
        return (out, lg) if want_logits else out

    def set_next_token(self, slot, tok):
        self._check(lib().llmc_seq_set_next_token(self.h, slot, int(tok)))

    def get_next_token(self, slot):
        t = C.c_int32(-1)
        self._check(lib().llmc_seq_get_next_token(self.h, slot, C.byref(t)))
        return t.value

    def stats(self):
        st = Stats()
        self._check(lib().llmc_engine_stats(self.h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in Stats._fields_}

    def debug_tensor(self, layer, kind):
        n = lib().llmc_debug_tensor(self.h, layer, kind, None, 0)
        if n < 0:
            raise LlmcError(int(n), "llmc_debug_tensor")
        out = np.empty(n, np.uint16)
        n2 = lib().llmc_debug_tensor(self.h, layer, kind, out.ctypes.data, n)
        if n2 != n:
            raise LlmcError(int(n2), "llmc_debug_tensor")
        return out

    def debug_kv(self, slot, layer, pos, kd, fp32=True):
        k = np.empty(kd, np.float32 if fp32 else np.uint16)
        v = np.empty(kd, np.float32 if fp32 else np.uint16)
        rc = lib().llmc_debug_kv(self.h, slot, layer, pos, k.ctypes.data, v.ctypes.data)
        if rc != OK:
            raise LlmcError(rc, "llmc_debug_kv")
        return k, v

    def generate(self, prompt, n_new, chunk=32, slot=None):
        """prefill + greedy decode of n_new tokens (the provider's inner loop); returns int32[n_new]."""
        own = slot is None
        if own:
            slot = self.seq_alloc()
        try:
            first = self.prefill(slot, prompt)
            toks = [first]
            left = n_new - 1
            while left > 0:
                k = min(chunk, left)
                toks.extend(self.decode([slot], k)[0].tolist())
                left -= k
            return np.asarray(toks[:n_new], np.int32)
        finally:
            if own:
                self.seq_free(slot)


# ------------------------------------------------------------------ kernel-level hooks
def _ops_check(rc):
    if rc != OK:
        raise LlmcError(rc, lib().llmc_ops_last_error().decode())


def _c(a, dt):
    return np.ascontiguousarray(a, dt)


def op_gemv(W, x, gain=None, eps=1e-5, y_in=None, device=0):
    W = _c(W, np.uint16); x = _c(x, np.float32)
    N, K = W.shape
    B = x.shape[0]
    y = _c(y_in, np.float32).copy() if y_in is not None else np.zeros((B, N), np.float32)
    g = _c(gain, np.uint16) if gain is not None else None
    _ops_check(lib().llmc_op_gemv(device, N, K, B, W.ctypes.data, x.ctypes.data, int(g is not None), _ptr(g), eps,
                                  int(y_in is not None), y.ctypes.data))
    return y


def op_gemv_swiglu(Wg, Wu, x, gain, eps=1e-5, device=0):
    Wg = _c(Wg, np.uint16); Wu = _c(Wu, np.uint16); x = _c(x, np.float32); g = _c(gain, np.uint16)
    I, K = Wg.shape
    B = x.shape[0]
    act = np.zeros((B, I), np.float32)
    _ops_check(lib().llmc_op_gemv_swiglu(device, I, K, B, Wg.ctypes.data, Wu.ctypes.data, x.ctypes.data, g.ctypes.data,
                                         eps, act.ctypes.data))
    return act


def op_lmhead(W, x, gain, eps=1e-5, want_logits=True, device=0):
    W = _c(W, np.uint16); x = _c(x, np.float32); g = _c(gain, np.uint16)
    V, K = W.shape
    B = x.shape[0]
    lg = np.zeros((B, V), np.float32) if want_logits else None
    tok = np.zeros(B, np.int32)
    _ops_check(lib().llmc_op_lmhead(device, V, K, B, W.ctypes.data, x.ctypes.data, g.ctypes.data, eps, _ptr(lg),
                                    tok.ctypes.data))
    return tok, lg


def op_attn_decode(q, k, v, ctx, q_heads, kv_heads, head_dim, page_size=64, nsplit=0, device=0):
    """k, v: uint16 arrays (bf16 bits) or float32 arrays select the KV element type."""
    kv_fp32 = int(np.asarray(k).dtype == np.float32)
    dt = np.float32 if kv_fp32 else np.uint16
    q = _c(q, np.float32); k = _c(k, dt); v = _c(v, dt); ctx = _c(ctx, np.int32)
    B, max_ctx = k.shape[0], k.shape[1]
    out = np.zeros_like(q)
    _ops_check(lib().llmc_op_attn_decode(device, B, q_heads, kv_heads, head_dim, page_size, nsplit, max_ctx, kv_fp32,
                                         ctx.ctypes.data, q.ctypes.data, k.ctypes.data, v.ctypes.data, out.ctypes.data))
    return out


def op_gemm(A, W, epi=1, impl=0, out_in=None, A_lo=None, want_lo=False, device=0):
    A = _c(A, np.uint16); W = _c(W, np.uint16)
    Al = _c(A_lo, np.uint16) if A_lo is not None else None
    T, K = A.shape
    N = W.shape[0]
    lo = None
    if epi == 0:
        out = np.zeros((T, N), np.uint16)
    elif epi == 1:
        out = np.zeros((T, N), np.float32)
    elif epi == 2:
        out = _c(out_in, np.float32).copy()
    else:
        out = np.zeros((T, N // 2), np.uint16)
        lo = np.zeros((T, N // 2), np.uint16) if want_lo else None
    _ops_check(lib().llmc_op_gemm(device, impl, epi, T, N, K, A.ctypes.data, _ptr(Al), W.ctypes.data, out.ctypes.data,
                                  _ptr(lo)))
    return (out, lo) if want_lo else out


def op_gemm_bench(T, N, K, epi=1, split=0, impl=0, iters=20, device=0):
    ms = C.c_float(0)
    _ops_check(lib().llmc_op_gemm_bench(device, impl, epi, split, T, N, K, iters, C.byref(ms)))
    return ms.value


def op_rmsnorm_rows(x, gain, eps=1e-5, split=False, device=0):
    x = _c(x, np.float32); g = _c(gain, np.uint16)
    T, H = x.shape
    out = np.zeros((T, H), np.uint16)
    lo = np.zeros((T, H), np.uint16) if split else None
    _ops_check(lib().llmc_op_rmsnorm_rows(device, T, H, x.ctypes.data, g.ctypes.data, eps, out.ctypes.data, _ptr(lo)))
    return (out, lo) if split else out


def op_prefill_attention(qkv, k_past, v_past, q_heads, kv_heads, head_dim, page_size=64, rope_theta=500000.0,
                         kv_fp32=True, split=True, device=0, impl=0):
    """Returns (out_hi, out_lo | None, q fp32, k_new, v_new); k/v arrays are float32 or uint16 (bf16 bits)."""
    dt = np.float32 if kv_fp32 else np.uint16
    qkv = _c(qkv, np.float32)
    T = qkv.shape[0]
    p0 = 0 if k_past is None else k_past.shape[0]
    kp = _c(k_past, dt) if p0 else None
    vp = _c(v_past, dt) if p0 else None
    QD, KD = q_heads * head_dim, kv_heads * head_dim
    out = np.zeros((T, QD), np.uint16)
    lo = np.zeros((T, QD), np.uint16) if split else None
    qo = np.zeros((T, QD), np.float32)
    kn = np.zeros((T, KD), dt); vn = np.zeros((T, KD), dt)
    _ops_check(lib().llmc_op_prefill_attention(device, T, p0, q_heads, kv_heads, head_dim, page_size, rope_theta,
                                               int(kv_fp32), qkv.ctypes.data, _ptr(kp), _ptr(vp), out.ctypes.data,
                                               _ptr(lo), qo.ctypes.data, kn.ctypes.data, vn.ctypes.data, impl))
    return out, lo, qo, kn, vn
