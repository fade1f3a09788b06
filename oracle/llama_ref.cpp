// oracle/llama_ref.cpp -- CPU fp32 ORACLE for the local provider's arithmetic.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under llm-consensus_b200/ may include, link or
// call this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs use it, and only as the checker / the timed CPU baseline.
//
// PARITY: UNPINNED BY THE REFERENCE.  johnayoung/llm-consensus contains no local
// inference arithmetic at all (every provider is a remote HTTPS call:
// /root/reference/internal/provider/openai.go:141-210, anthropic.go:132-202,
// google.go:136-208), so there is no reference file this forward pass could follow
// and no golden vector for it.  BASELINE.json's north_star defines the oracle
// instead: "a pure-Go fp32 CPU forward of the same random-init architecture".  Go is
// not installed in this image, so the oracle is written in C++17 (OpenMP) -- the
// runnable form -- and it is pinned two other ways (tests/test_oracle_pin.py):
//   * against an independent numpy restatement (oracle/llama_np.py), and
//   * against HuggingFace transformers' LlamaForCausalLM on CPU fp32 loaded with the
//     same weights (fixtures committed under tests/golden/, generator script beside).
// The boundary the forward sits behind is provider.Provider.QueryStream
// (/root/reference/internal/provider/provider.go:13-21).
//
// Architecture (Llama-3 family, as fixed in SURVEY.md section 7 step 1):
//   x = E[tok]
//   per layer: h = rmsnorm(x, g1); q,k,v = Wq h, Wk h, Wv h; rope(q,k) rotate-half,
//              theta = 5e5; causal GQA attention with fp32 softmax; x += Wo a;
//              h = rmsnorm(x, g2); x += Wd (silu(Wg h) * (Wu h))
//   logits = LMHead rmsnorm(x, gf)     (LMHead = E when tied); greedy = lowest-index argmax
// Weights are the bf16-rounded values of a counter-based RNG (see rng_* below),
// widened to fp32; all activations and accumulations are fp32.
//
// Build: see oracle/Makefile (g++ -O3 -fopenmp -shared -fPIC).
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

typedef struct {
  int32_t hidden, layers, q_heads, kv_heads, head_dim, inter, vocab, tied;
  float rope_theta, rms_eps, init_std;
  uint64_t seed;
  int32_t max_ctx;
  // RoPE frequency scaling as in Hugging Face's rope_type "llama3" (transformers modeling_rope_utils.py,
  // _compute_llama3_parameters): 0 = none (rope_type "default"), 1 = llama3 with the four parameters below
  int32_t rope_scaling;
  float rope_factor, rope_low_freq_factor, rope_high_freq_factor;
  int32_t rope_original_ctx;
} oref_config;

}  // extern "C"

// ------------------------------------------------------------------ RNG spec
// The weight definition shared (by specification, not by code) with the CUDA init
// kernels in llm-consensus_b200/csrc/init.cu:
//   key   = splitmix64_mix(seed ^ (GOLDEN * (tensor_id + 1)))
//   z     = splitmix64_mix(key + GOLDEN * (index + 1))
//   u     = (int32(z >> 40) - 2^23) * 2^-23              in [-1, 1), exact in fp32
//   W[i]  = bf16_rne(u * amp),  amp = float(init_std * sqrt(3))      (matrices)
//   g[i]  = bf16_rne(1 + 0.1f * u)   (two separately rounded fp32 ops) (norm gains)
// tensor_id = (layer + 1) * 16 + kind; globals use layer = -1.
static const uint64_t GOLDEN = 0x9E3779B97F4A7C15ull;
static inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static inline uint64_t rng_key(uint64_t seed, uint32_t tensor_id) {
  return mix64(seed ^ (GOLDEN * (uint64_t)(tensor_id + 1)));
}
static inline float rng_unit(uint64_t key, uint64_t idx) {
  uint64_t z = mix64(key + GOLDEN * (idx + 1));
  int32_t v = (int32_t)(z >> 40) - (1 << 23);
  return (float)v * (1.0f / 8388608.0f);
}
static inline uint16_t bf16_rne(float f) {
  uint32_t b;
  memcpy(&b, &f, 4);
  if ((b & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((b >> 16) | 0x40);  // NaN
  b += 0x7fffu + ((b >> 16) & 1u);
  return (uint16_t)(b >> 16);
}
static inline float bf16_to_f32(uint16_t h) {
  uint32_t b = (uint32_t)h << 16;
  float f;
  memcpy(&f, &b, 4);
  return f;
}

enum { K_EMB = 0, K_LMHEAD = 1, K_FINAL_NORM = 2 };                               // globals
enum { K_ATTN_NORM = 0, K_WQ, K_WK, K_WV, K_WO, K_MLP_NORM, K_WG, K_WU, K_WD };  // per layer
static inline uint32_t tensor_id(int layer, int kind) { return (uint32_t)((layer + 1) * 16 + kind); }

static void fill_matrix(std::vector<uint16_t>& w, size_t n, uint64_t seed, uint32_t tid, float amp) {
  w.resize(n);
  uint64_t key = rng_key(seed, tid);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i) w[i] = bf16_rne(rng_unit(key, (uint64_t)i) * amp);
}
static void fill_gain(std::vector<uint16_t>& w, size_t n, uint64_t seed, uint32_t tid) {
  w.resize(n);
  uint64_t key = rng_key(seed, tid);
  for (size_t i = 0; i < n; ++i) {
    volatile float p = 0.1f * rng_unit(key, i);  // volatile: forbid fma contraction
    w[i] = bf16_rne(1.0f + p);
  }
}

struct Layer {
  std::vector<uint16_t> g1, wq, wk, wv, wo, g2, wg, wu, wd;
};
struct Model {
  oref_config c;
  std::vector<uint16_t> emb, lmhead, gf;
  std::vector<Layer> L;
  // dense KV cache, fp32: [layer][pos][kv_heads*head_dim]
  std::vector<float> kc, vc;
  std::vector<float> rope_cos, rope_sin;  // [max_ctx][head_dim/2]
  int len = 0;
  // DIAGNOSTIC ONLY (never used by parity tests): emulate bf16 rounding of activations at the
  // sites where the GPU prefill stores bf16, to attribute the GPU-vs-fp32 logit error.
  //   1: normalised GEMM inputs  2: q  4: k,v (cache)  8: attention output  16: SwiGLU output
  int round_mask = 0;
};

// Y[n,N] (+)= X[n,K] * W[N,K]^T, W bf16 widened to fp32, fp32 accumulation.
static void linear(const float* X, int n, int K, const uint16_t* W, int N, float* Y, int ldy) {
  const int TB = 32;
  for (int t0 = 0; t0 < n; t0 += TB) {
    int tn = std::min(TB, n - t0);
#pragma omp parallel
    {
      std::vector<float> wrow(K);
#pragma omp for schedule(static)
      for (int j = 0; j < N; ++j) {
        const uint16_t* wr = W + (size_t)j * K;
        for (int k = 0; k < K; ++k) wrow[k] = bf16_to_f32(wr[k]);
        for (int t = 0; t < tn; ++t) {
          const float* x = X + (size_t)(t0 + t) * K;
          float acc = 0.f;
#pragma omp simd reduction(+ : acc)
          for (int k = 0; k < K; ++k) acc += x[k] * wrow[k];
          Y[(size_t)(t0 + t) * ldy + j] = acc;
        }
      }
    }
  }
}

static inline void round_bf16_inplace(float* v, size_t n) {
  for (size_t i = 0; i < n; ++i) v[i] = bf16_to_f32(bf16_rne(v[i]));
}

static void rmsnorm(const float* x, const uint16_t* g, int H, float eps, float* y) {
  float ss = 0.f;
  for (int i = 0; i < H; ++i) ss += x[i] * x[i];
  float r = 1.0f / sqrtf(ss / (float)H + eps);
  for (int i = 0; i < H; ++i) y[i] = x[i] * r * bf16_to_f32(g[i]);
}

extern "C" {

void* oref_create(const oref_config* cfg) {
  Model* m = new Model();
  m->c = *cfg;
  const oref_config& c = m->c;
  const int H = c.hidden, D = c.head_dim, QD = c.q_heads * D, KD = c.kv_heads * D, I = c.inter;
  float amp = (float)((double)c.init_std * 1.7320508075688772);
  fill_matrix(m->emb, (size_t)c.vocab * H, c.seed, tensor_id(-1, K_EMB), amp);
  if (!c.tied) fill_matrix(m->lmhead, (size_t)c.vocab * H, c.seed, tensor_id(-1, K_LMHEAD), amp);
  fill_gain(m->gf, H, c.seed, tensor_id(-1, K_FINAL_NORM));
  m->L.resize(c.layers);
  for (int l = 0; l < c.layers; ++l) {
    Layer& y = m->L[l];
    fill_gain(y.g1, H, c.seed, tensor_id(l, K_ATTN_NORM));
    fill_matrix(y.wq, (size_t)QD * H, c.seed, tensor_id(l, K_WQ), amp);
    fill_matrix(y.wk, (size_t)KD * H, c.seed, tensor_id(l, K_WK), amp);
    fill_matrix(y.wv, (size_t)KD * H, c.seed, tensor_id(l, K_WV), amp);
    fill_matrix(y.wo, (size_t)H * QD, c.seed, tensor_id(l, K_WO), amp);
    fill_gain(y.g2, H, c.seed, tensor_id(l, K_MLP_NORM));
    fill_matrix(y.wg, (size_t)I * H, c.seed, tensor_id(l, K_WG), amp);
    fill_matrix(y.wu, (size_t)I * H, c.seed, tensor_id(l, K_WU), amp);
    fill_matrix(y.wd, (size_t)H * I, c.seed, tensor_id(l, K_WD), amp);
  }
  m->kc.assign((size_t)c.layers * c.max_ctx * KD, 0.f);
  m->vc.assign((size_t)c.layers * c.max_ctx * KD, 0.f);
  m->rope_cos.resize((size_t)c.max_ctx * (D / 2));
  m->rope_sin.resize((size_t)c.max_ctx * (D / 2));
  for (int p = 0; p < c.max_ctx; ++p)
    for (int i = 0; i < D / 2; ++i) {
      double inv = pow((double)c.rope_theta, -2.0 * i / (double)D);
      if (c.rope_scaling == 1) {
        // wavelengths longer than original_ctx / low_freq_factor are stretched by `factor`, those shorter than
        // original_ctx / high_freq_factor are kept, the band between is interpolated
        const double two_pi = 6.283185307179586476925286766559;
        const double wavelen = two_pi / inv, oc = (double)c.rope_original_ctx;
        const double lo_w = oc / (double)c.rope_low_freq_factor, hi_w = oc / (double)c.rope_high_freq_factor;
        if (wavelen > lo_w) inv = inv / (double)c.rope_factor;
        else if (!(wavelen < hi_w)) {
          const double smooth = (oc / wavelen - (double)c.rope_low_freq_factor) /
                                ((double)c.rope_high_freq_factor - (double)c.rope_low_freq_factor);
          inv = (1.0 - smooth) * inv / (double)c.rope_factor + smooth * inv;
        }
      }
      double a = (double)p * inv;
      m->rope_cos[(size_t)p * (D / 2) + i] = (float)cos(a);
      m->rope_sin[(size_t)p * (D / 2) + i] = (float)sin(a);
    }
  return m;
}

void oref_destroy(void* h) { delete (Model*)h; }
void oref_reset(void* h) { ((Model*)h)->len = 0; }
void oref_set_round_mask(void* h, int mask) { ((Model*)h)->round_mask = mask; }
int oref_len(void* h) { return ((Model*)h)->len; }
// Truncate the context to `len` tokens (cache entries beyond are simply overwritten later).
void oref_truncate(void* h, int len) {
  Model* m = (Model*)h;
  if (len >= 0 && len <= m->len) m->len = len;
}

// Copy a weight tensor (bf16 bits) out, for pinning the GPU initialiser.  layer=-1: globals.
// Returns element count, or -1.
int64_t oref_tensor(void* h, int layer, int kind, uint16_t* out, int64_t cap) {
  Model* m = (Model*)h;
  const std::vector<uint16_t>* v = nullptr;
  if (layer < 0) {
    if (kind == K_EMB) v = &m->emb;
    else if (kind == K_LMHEAD) v = m->c.tied ? &m->emb : &m->lmhead;
    else if (kind == K_FINAL_NORM) v = &m->gf;
  } else if (layer < m->c.layers) {
    Layer& y = m->L[layer];
    const std::vector<uint16_t>* t[] = {&y.g1, &y.wq, &y.wk, &y.wv, &y.wo, &y.g2, &y.wg, &y.wu, &y.wd};
    if (kind >= 0 && kind < 9) v = t[kind];
  }
  if (!v) return -1;
  int64_t n = (int64_t)v->size();
  if (out && cap >= n) memcpy(out, v->data(), (size_t)n * 2);
  return n;
}

// Append n tokens to the context and run them through the model.
//   logits_mode 0: no logits; 1: last position only -> logits[V]; 2: all n -> logits[n][V]
// Returns 0, or -1 if the context would overflow.
int oref_forward(void* h, const int32_t* tokens, int n, int logits_mode, float* logits) {
  Model* m = (Model*)h;
  const oref_config& c = m->c;
  const int H = c.hidden, D = c.head_dim, NQ = c.q_heads, NKV = c.kv_heads;
  const int QD = NQ * D, KD = NKV * D, I = c.inter, V = c.vocab, G = NQ / NKV;
  if (n <= 0) return 0;
  if (m->len + n > c.max_ctx) return -1;
  for (int t = 0; t < n; ++t)
    if (tokens[t] < 0 || tokens[t] >= V) return -2;
  const int p0 = m->len;
  std::vector<float> x((size_t)n * H), hn((size_t)n * H), q((size_t)n * QD), att((size_t)n * QD);
  std::vector<float> kv((size_t)n * KD), gate((size_t)n * I), up((size_t)n * I), tmp((size_t)n * H);
  for (int t = 0; t < n; ++t)
    for (int i = 0; i < H; ++i) x[(size_t)t * H + i] = bf16_to_f32(m->emb[(size_t)tokens[t] * H + i]);
  const float scale = 1.0f / sqrtf((float)D);
  for (int l = 0; l < c.layers; ++l) {
    Layer& y = m->L[l];
    float* kc = m->kc.data() + (size_t)l * c.max_ctx * KD;
    float* vc = m->vc.data() + (size_t)l * c.max_ctx * KD;
#pragma omp parallel for schedule(static)
    for (int t = 0; t < n; ++t) rmsnorm(&x[(size_t)t * H], y.g1.data(), H, c.rms_eps, &hn[(size_t)t * H]);
    if (m->round_mask & 1) round_bf16_inplace(hn.data(), (size_t)n * H);
    linear(hn.data(), n, H, y.wq.data(), QD, q.data(), QD);
    linear(hn.data(), n, H, y.wk.data(), KD, kc + (size_t)p0 * KD, KD);
    linear(hn.data(), n, H, y.wv.data(), KD, vc + (size_t)p0 * KD, KD);
    // RoPE (rotate-half) on q and the new k rows
#pragma omp parallel for schedule(static)
    for (int t = 0; t < n; ++t) {
      const float* cs = &m->rope_cos[(size_t)(p0 + t) * (D / 2)];
      const float* sn = &m->rope_sin[(size_t)(p0 + t) * (D / 2)];
      for (int hh = 0; hh < NQ + NKV; ++hh) {
        float* v = hh < NQ ? &q[(size_t)t * QD + hh * D] : kc + (size_t)(p0 + t) * KD + (hh - NQ) * D;
        for (int i = 0; i < D / 2; ++i) {
          float a = v[i], b = v[i + D / 2];
          v[i] = a * cs[i] - b * sn[i];
          v[i + D / 2] = b * cs[i] + a * sn[i];
        }
      }
    }
    if (m->round_mask & 2) round_bf16_inplace(q.data(), (size_t)n * QD);
    if (m->round_mask & 4) {
      round_bf16_inplace(kc + (size_t)p0 * KD, (size_t)n * KD);
      round_bf16_inplace(vc + (size_t)p0 * KD, (size_t)n * KD);
    }
    // causal GQA attention, fp32 softmax
#pragma omp parallel for schedule(dynamic, 4) collapse(2)
    for (int t = 0; t < n; ++t)
      for (int hh = 0; hh < NQ; ++hh) {
        const int ctx = p0 + t + 1, kvh = hh / G;
        const float* qv = &q[(size_t)t * QD + hh * D];
        std::vector<float> s(ctx);
        float mx = -INFINITY;
        for (int p = 0; p < ctx; ++p) {
          const float* kr = kc + (size_t)p * KD + kvh * D;
          float acc = 0.f;
          for (int i = 0; i < D; ++i) acc += qv[i] * kr[i];
          s[p] = acc * scale;
          mx = std::max(mx, s[p]);
        }
        float den = 0.f;
        for (int p = 0; p < ctx; ++p) {
          s[p] = expf(s[p] - mx);
          den += s[p];
        }
        float* o = &att[(size_t)t * QD + hh * D];
        for (int i = 0; i < D; ++i) o[i] = 0.f;
        for (int p = 0; p < ctx; ++p) {
          const float* vr = vc + (size_t)p * KD + kvh * D;
          float w = s[p] / den;
          for (int i = 0; i < D; ++i) o[i] += w * vr[i];
        }
      }
    if (m->round_mask & 8) round_bf16_inplace(att.data(), (size_t)n * QD);
    linear(att.data(), n, QD, y.wo.data(), H, tmp.data(), H);
    for (size_t i = 0; i < (size_t)n * H; ++i) x[i] += tmp[i];
#pragma omp parallel for schedule(static)
    for (int t = 0; t < n; ++t) rmsnorm(&x[(size_t)t * H], y.g2.data(), H, c.rms_eps, &hn[(size_t)t * H]);
    if (m->round_mask & 1) round_bf16_inplace(hn.data(), (size_t)n * H);
    linear(hn.data(), n, H, y.wg.data(), I, gate.data(), I);
    linear(hn.data(), n, H, y.wu.data(), I, up.data(), I);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n * I; ++i) {
      float g = gate[i];
      gate[i] = g / (1.0f + expf(-g)) * up[i];
    }
    if (m->round_mask & 16) round_bf16_inplace(gate.data(), (size_t)n * I);
    linear(gate.data(), n, I, y.wd.data(), H, tmp.data(), H);
    for (size_t i = 0; i < (size_t)n * H; ++i) x[i] += tmp[i];
  }
  m->len += n;
  if (logits_mode == 0 || !logits) return 0;
  const uint16_t* head = c.tied ? m->emb.data() : m->lmhead.data();
  if (logits_mode == 1) {
    rmsnorm(&x[(size_t)(n - 1) * H], m->gf.data(), H, c.rms_eps, hn.data());
    linear(hn.data(), 1, H, head, V, logits, V);
  } else {
    for (int t = 0; t < n; ++t) rmsnorm(&x[(size_t)t * H], m->gf.data(), H, c.rms_eps, &hn[(size_t)t * H]);
    linear(hn.data(), n, H, head, V, logits, V);
  }
  return 0;
}

// Greedy argmax, lowest index wins ties.
int32_t oref_argmax(const float* v, int n) {
  int32_t best = 0;
  for (int i = 1; i < n; ++i)
    if (v[i] > v[best]) best = i;
  return best;
}

// prefill `prompt` then greedy-decode n_new tokens.  out_tokens[n_new]; if logits != NULL it
// receives [n_new][V] (the logits each generated token was the argmax of).
int oref_generate(void* h, const int32_t* prompt, int n_prompt, int n_new, int32_t* out_tokens, float* logits) {
  Model* m = (Model*)h;
  const int V = m->c.vocab;
  std::vector<float> lg(V);
  int rc = oref_forward(h, prompt, n_prompt, 1, lg.data());
  if (rc) return rc;
  for (int s = 0; s < n_new; ++s) {
    int32_t t = oref_argmax(lg.data(), V);
    out_tokens[s] = t;
    if (logits) memcpy(logits + (size_t)s * V, lg.data(), (size_t)V * 4);
    if (s + 1 < n_new) {
      rc = oref_forward(h, &t, 1, 1, lg.data());
      if (rc) return rc;
    }
  }
  return 0;
}

int oref_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
