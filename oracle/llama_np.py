"""Independent numpy restatement of the oracle's arithmetic (weights RNG + forward), plus the
per-operator pieces the kernel-level parity tests use.

TEST INFRASTRUCTURE ONLY (see oracle/ref.py).  It exists so that the C++ oracle
(oracle/llama_ref.cpp) is pinned by a second, separately written implementation; the
reference repository offers no arithmetic to pin it against (parity unpinned by the
reference -- its providers are remote HTTP calls, /root/reference/internal/provider/*.go).
"""
import numpy as np

GOLDEN = np.uint64(0x9E3779B97F4A7C15)
M1 = np.uint64(0xBF58476D1CE4E5B9)
M2 = np.uint64(0x94D049BB133111EB)


def _mix64(z):
    z = np.asarray(z, np.uint64)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * M1
        z = (z ^ (z >> np.uint64(27))) * M2
        return z ^ (z >> np.uint64(31))


def tensor_id(layer, kind):
    return (layer + 1) * 16 + kind


def rng_unit(seed, tid, n, start=0):
    """u in [-1,1) for indices start..start+n-1 of tensor `tid`."""
    with np.errstate(over="ignore"):
        key = _mix64(np.uint64(seed) ^ (GOLDEN * np.uint64(tid + 1)))
        idx = np.arange(start, start + n, dtype=np.uint64)
        z = _mix64(key + GOLDEN * (idx + np.uint64(1)))
    v = (z >> np.uint64(40)).astype(np.int64) - (1 << 23)
    return v.astype(np.float32) * np.float32(1.0 / 8388608.0)


def bf16_rne_bits(f):
    b = np.ascontiguousarray(f, np.float32).view(np.uint32).astype(np.uint64)
    b = b + np.uint64(0x7FFF) + ((b >> np.uint64(16)) & np.uint64(1))
    return (b >> np.uint64(16)).astype(np.uint16)


def bf16_to_f32(u16):
    return (np.asarray(u16, np.uint16).astype(np.uint32) << 16).view(np.float32)


def bf16_round(f):
    return bf16_to_f32(bf16_rne_bits(f))


def matrix(seed, tid, rows, cols, init_std=0.02):
    amp = np.float32(np.float64(np.float32(init_std)) * 1.7320508075688772)
    u = rng_unit(seed, tid, rows * cols)
    return bf16_round(u * amp).reshape(rows, cols)


def gain(seed, tid, n):
    u = rng_unit(seed, tid, n)
    p = (np.float32(0.1) * u).astype(np.float32)
    return bf16_round((np.float32(1.0) + p).astype(np.float32))


# ------------------------------------------------------------------ operators (fp32)
def rmsnorm(x, g, eps=1e-5):
    x = np.asarray(x, np.float32)
    ms = np.mean(x.astype(np.float32) ** 2, axis=-1, keepdims=True, dtype=np.float32)
    return (x * (np.float32(1.0) / np.sqrt(ms + np.float32(eps))) * g).astype(np.float32)


def rope_tables(max_ctx, head_dim, theta=500000.0):
    i = np.arange(head_dim // 2, dtype=np.float64)
    inv = np.power(np.float64(theta), -2.0 * i / head_dim)
    a = np.arange(max_ctx, dtype=np.float64)[:, None] * inv[None, :]
    return np.cos(a).astype(np.float32), np.sin(a).astype(np.float32)


def rope(v, pos, cos, sin):
    """v [..., heads, d]; pos scalar or [T] matching v's first axis; rotate-half."""
    d = v.shape[-1]
    a, b = v[..., : d // 2], v[..., d // 2:]
    c = cos[pos]
    s = sin[pos]
    if np.ndim(pos) == 1:
        c = c[:, None, :]
        s = s[:, None, :]
    return np.concatenate([a * c - b * s, b * c + a * s], axis=-1).astype(np.float32)


def attention(q, k, v):
    """q [nq, d] for one position; k,v [ctx, nkv, d]; GQA; returns [nq, d] (fp32 softmax)."""
    nq, d = q.shape
    nkv = k.shape[1]
    g = nq // nkv
    out = np.empty_like(q)
    for h in range(nq):
        s = (k[:, h // g, :] @ q[h]) * np.float32(1.0 / np.sqrt(d))
        p = np.exp(s - s.max())
        p = p / p.sum(dtype=np.float32)
        out[h] = p.astype(np.float32) @ v[:, h // g, :]
    return out.astype(np.float32)


def swiglu(gate, up):
    gate = np.asarray(gate, np.float32)
    return (gate / (np.float32(1.0) + np.exp(-gate)) * up).astype(np.float32)


class NumpyLlama:
    """Straight-line fp32 forward; slow, for tiny shapes only."""

    def __init__(self, shape, seed=0, max_ctx=256, rope_theta=500000.0, rms_eps=1e-5, init_std=0.02):
        self.s = dict(shape)
        s = self.s
        H, D = s["hidden"], s["head_dim"]
        QD, KD, I, V = s["q_heads"] * D, s["kv_heads"] * D, s["inter"], s["vocab"]
        self.eps = rms_eps
        self.emb = matrix(seed, tensor_id(-1, 0), V, H, init_std)
        self.head = self.emb if s["tied"] else matrix(seed, tensor_id(-1, 1), V, H, init_std)
        self.gf = gain(seed, tensor_id(-1, 2), H)
        self.layers = []
        for l in range(s["layers"]):
            t = lambda k: tensor_id(l, k)
            self.layers.append(dict(
                g1=gain(seed, t(0), H), wq=matrix(seed, t(1), QD, H, init_std), wk=matrix(seed, t(2), KD, H, init_std),
                wv=matrix(seed, t(3), KD, H, init_std), wo=matrix(seed, t(4), H, QD, init_std),
                g2=gain(seed, t(5), H), wg=matrix(seed, t(6), I, H, init_std), wu=matrix(seed, t(7), I, H, init_std),
                wd=matrix(seed, t(8), H, I, init_std)))
        self.cos, self.sin = rope_tables(max_ctx, D, rope_theta)

    def forward_all(self, tokens):
        """Full-sequence causal forward; returns logits [T, V]."""
        s = self.s
        D, NQ, NKV = s["head_dim"], s["q_heads"], s["kv_heads"]
        T = len(tokens)
        x = self.emb[np.asarray(tokens)].astype(np.float32)
        pos = np.arange(T)
        for L in self.layers:
            h = rmsnorm(x, L["g1"], self.eps)
            q = rope((h @ L["wq"].T).reshape(T, NQ, D), pos, self.cos, self.sin)
            k = rope((h @ L["wk"].T).reshape(T, NKV, D), pos, self.cos, self.sin)
            v = (h @ L["wv"].T).reshape(T, NKV, D)
            a = np.stack([attention(q[t], k[: t + 1], v[: t + 1]) for t in range(T)]).reshape(T, NQ * D)
            x = x + a @ L["wo"].T
            h = rmsnorm(x, L["g2"], self.eps)
            x = x + swiglu(h @ L["wg"].T, h @ L["wu"].T) @ L["wd"].T
        return (rmsnorm(x, self.gf, self.eps) @ self.head.T).astype(np.float32)
