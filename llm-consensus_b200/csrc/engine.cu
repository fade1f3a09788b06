// engine.cu -- the inference engine behind the C ABI in include/llmc.h: one model instance on
// one GPU (weights generated in HBM from a seed, paged bf16 KV pool, prefill and decode
// schedules over the kernels in this directory).
//
// It stands where a remote backend's HTTP round trip is in the reference:
// QueryStream in /root/reference/internal/provider/openai.go:141-210 sends the prompt and then
// scans SSE lines; a local provider calls llmc_prefill() and then loops over llmc_decode().
//
// HBM layout (per engine):
//   weights   one blob, bf16, every matrix row-major [out][in] (K-major: the same bytes feed
//             the decode GEMV stream and the TMA tiles of the prefill GEMM)
//               per layer: g1[H] | Wqkv[(nq+2nkv)d][H] | Wo[H][nq d] | g2[H] |
//                          Wgu[2I][H] (gate/up interleaved in blocks of 64 rows) | Wd[H][I]
//               global:    E[V][H] | LMHead[V][H] (= E when tied) | gf[H]
//   KV pool   K and V: [layer][page][kv_head][page_size][d], fp32 (default) or bf16; block table [slot][page]
//   decode    x[B][H], q[B][nq d], attn[B][nq d], act[B][I] fp32; split-attention scratch
//   prefill   x[T][H] fp32, qkv[T][(nq+2nkv)d] fp32, q[T][nq d] fp32; GEMM operands xn[T][H],
//             attn[T][nq d], act[T][I] as bf16 hi (+ lo) pairs, for T = prefill_chunk
#include <cuda_runtime.h>
#include <chrono>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/llmc.h"
#include "engine_internal.h"

using namespace llmc;

namespace {
thread_local std::string g_create_error;
}  // namespace

namespace {

int fail(llmc_engine* e, int code, const std::string& msg) {
  if (e) e->last_error = msg;
  else g_create_error = msg;
  return code;
}

#define E_TRY(expr)                                                                                        \
  do {                                                                                                     \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess)                                                                                 \
      return fail(e, _e == cudaErrorMemoryAllocation ? LLMC_ERR_OOM : LLMC_ERR_CUDA,                       \
                  std::string(#expr) + ": " + cudaGetErrorString(_e) + " (engine.cu:" + std::to_string(__LINE__) + ")"); \
  } while (0)

template <typename T>
int dalloc(llmc_engine* e, T** p, size_t n) {
  void* v = nullptr;
  E_TRY(cudaMalloc(&v, n * sizeof(T) + 256));
  e->allocs.push_back(v);
  *p = reinterpret_cast<T*>(v);
  return LLMC_OK;
}

int apply_shape(llmc_config& c) {
  if (c.shape == LLMC_SHAPE_LLAMA_1B) {
    c.hidden = 2048; c.layers = 16; c.q_heads = 32; c.kv_heads = 8; c.head_dim = 64; c.inter = 8192;
    c.vocab = 128256; c.tied = 1;
  } else if (c.shape == LLMC_SHAPE_LLAMA_8B) {
    c.hidden = 4096; c.layers = 32; c.q_heads = 32; c.kv_heads = 8; c.head_dim = 128; c.inter = 14336;
    c.vocab = 128256; c.tied = 0;
  } else if (c.shape != LLMC_SHAPE_CUSTOM) {
    return LLMC_ERR_INVALID;
  }
  return LLMC_OK;
}

std::string validate(const llmc_config& c) {
  auto bad = [](const char* s) { return std::string("unsupported shape: ") + s; };
  if (c.hidden < 256 || c.hidden % 256) return bad("hidden must be a positive multiple of 256");
  if (c.layers < 1) return bad("layers < 1");
  if (c.head_dim != 32 && c.head_dim != 64 && c.head_dim != 128) return bad("head_dim must be 32, 64 or 128");
  if (c.kv_heads < 1 || c.q_heads % c.kv_heads) return bad("q_heads must be a multiple of kv_heads");
  int g = c.q_heads / c.kv_heads;
  if (g != 1 && g != 2 && g != 4 && g != 8) return bad("q_heads/kv_heads must be 1, 2, 4 or 8");
  if ((c.q_heads * c.head_dim) % 256) return bad("q_heads*head_dim must be a multiple of 256");
  if (((c.q_heads + 2 * c.kv_heads) * c.head_dim) % 64) return bad("(q_heads+2*kv_heads)*head_dim must be a multiple of 64");
  if (c.inter < 256 || c.inter % 256) return bad("inter must be a positive multiple of 256");
  if (c.vocab < 4 || c.vocab % 4) return bad("vocab must be a positive multiple of 4");
  if (c.page_size < 32 || c.page_size % 32) return "page_size must be a positive multiple of 32";
  if (c.max_slots < 1 || c.max_slots > 1024) return "max_slots out of range [1,1024]";
  if (c.max_ctx < 32) return "max_ctx < 32";
  if (c.prefill_chunk < 32) return "prefill_chunk < 32";
  if (c.prefill_impl != 0 && c.prefill_impl != 1) return "prefill_impl must be 0 or 1";
  return "";
}

int init_weights(llmc_engine* e, const char* weights_path) {
  const ModelDims& d = e->d;
  const int H = d.hidden, QD = d.qd(), KD = d.kd(), I = d.inter, V = d.vocab;
  const size_t per_layer = (size_t)2 * H + (size_t)(QD + 2 * KD) * H + (size_t)H * QD + (size_t)2 * I * H + (size_t)H * I;
  const size_t global = (size_t)V * H * (d.tied ? 1 : 2) + H;
  e->wblob_elems = per_layer * d.layers + global;
  if (int rc = dalloc(e, &e->wblob, e->wblob_elems)) return rc;
  const float amp = (float)((double)d.init_std * 1.7320508075688772);
  uint16_t* p = e->wblob;
  auto take = [&](size_t n) { uint16_t* r = p; p += n; return r; };
  cudaStream_t s = e->stream;
  e->L.resize(d.layers);
  for (int l = 0; l < d.layers; ++l) {
    LayerW& w = e->L[l];
    // matrices first so that every matrix start stays 16-byte aligned for any H % 8 == 0
    w.wqkv = take((size_t)(QD + 2 * KD) * H);
    w.wo = take((size_t)H * QD);
    w.wgu = take((size_t)2 * I * H);
    w.wd = take((size_t)H * I);
    w.g1 = take(H);
    w.g2 = take(H);
    if (weights_path) continue;   // layout only; filled from the checkpoint below
    E_TRY(launch_init_gain(w.g1, H, e->seed, tensor_id(l, K_ATTN_NORM), s));
    E_TRY(launch_init_matrix(w.wqkv, QD, H, H, e->seed, tensor_id(l, K_WQ), amp, 0, 0, s));
    E_TRY(launch_init_matrix(w.wqkv, KD, H, H, e->seed, tensor_id(l, K_WK), amp, 0, QD, s));
    E_TRY(launch_init_matrix(w.wqkv, KD, H, H, e->seed, tensor_id(l, K_WV), amp, 0, QD + KD, s));
    E_TRY(launch_init_matrix(w.wo, H, QD, QD, e->seed, tensor_id(l, K_WO), amp, 0, 0, s));
    E_TRY(launch_init_gain(w.g2, H, e->seed, tensor_id(l, K_MLP_NORM), s));
    E_TRY(launch_init_matrix(w.wgu, I, H, H, e->seed, tensor_id(l, K_WG), amp, 1, 0, s));
    E_TRY(launch_init_matrix(w.wgu, I, H, H, e->seed, tensor_id(l, K_WU), amp, 2, 0, s));
    E_TRY(launch_init_matrix(w.wd, H, I, I, e->seed, tensor_id(l, K_WD), amp, 0, 0, s));
  }
  e->emb = take((size_t)V * H);
  if (!weights_path) E_TRY(launch_init_matrix(e->emb, V, H, H, e->seed, tensor_id(-1, K_EMB), amp, 0, 0, s));
  if (d.tied) {
    e->lmhead = e->emb;
  } else {
    e->lmhead = take((size_t)V * H);
    if (!weights_path) E_TRY(launch_init_matrix(e->lmhead, V, H, H, e->seed, tensor_id(-1, K_LMHEAD), amp, 0, 0, s));
  }
  e->gf = take(H);
  if (!weights_path) E_TRY(launch_init_gain(e->gf, H, e->seed, tensor_id(-1, K_FINAL_NORM), s));
  if (weights_path) {
    std::string err;
    if (int rc = load_checkpoint(e, weights_path, &err)) return fail(e, rc, "weights_path: " + err);
  } else {
    e->stats.kernel_launches += (uint64_t)d.layers * 9 + 3;
  }
  e->stats.weight_bytes = 2.0 * ((double)per_layer * d.layers + (double)V * H + H);
  e->stats.kv_bytes_per_token = 2.0 * e->kv_elem * KD * d.layers;
  return LLMC_OK;
}

int ensure_pages(llmc_engine* e, int slot, int need_tokens) {
  if (need_tokens > e->max_ctx)
    return fail(e, LLMC_ERR_LIMIT, "sequence would exceed max_ctx (" + std::to_string(e->max_ctx) + " tokens)");
  const int need = (need_tokens + e->page_size - 1) / e->page_size;
  int have = e->slot_npages[slot];
  if (need <= have) return LLMC_OK;
  if ((int)e->free_pages.size() < need - have) return fail(e, LLMC_ERR_OOM, "KV page pool exhausted");
  int* row = &e->h_block_table[(size_t)slot * e->pages_per_seq];
  for (int i = have; i < need; ++i) {
    row[i] = e->free_pages.back();
    e->free_pages.pop_back();
    e->page_ref[row[i]] = 1;
  }
  E_TRY(cudaMemcpyAsync(e->d_block_table + (size_t)slot * e->pages_per_seq + have, row + have,
                        (size_t)(need - have) * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  e->slot_npages[slot] = need;
  return LLMC_OK;
}

int gemm(llmc_engine* e, const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int T,
         int N, int K, int epi, bool decode = false) {
  // up to 32 rows the swapped-operand decode kernel wins (deeper weight pipeline); above, the prefill kernel + split-K
  if (e->prefill_impl == 0 && decode && T <= 32 && gemm_decode_tc_supported(T, N, K, epi))
    E_TRY(launch_gemm_decode_tc(A, Alo, W, out, out_lo, T, N, K, epi, e->stream, e->use_pdl));
  else if (e->prefill_impl == 0)
    E_TRY(launch_gemm_tcgen05(A, Alo, W, out, out_lo, T, N, K, epi, e->stream, decode, decode && e->use_pdl));
  else E_TRY(launch_gemm_simple(A, Alo, W, out, out_lo, T, N, K, epi, e->stream));
  e->stats.kernel_launches++;
  return LLMC_OK;
}

GemvArgs base_args(llmc_engine* e, int B) {
  GemvArgs a;
  memset(&a, 0, sizeof(a));
  const ModelDims& d = e->d;
  a.B = B;
  a.eps = d.rms_eps;
  a.slots = e->d_slots;
  a.next_token = e->d_next_token;
  a.rope_cos = e->rope_cos;
  a.rope_sin = e->rope_sin;
  a.seq_len = e->d_seq_len;
  a.block_table = e->d_block_table;
  a.bt_stride = e->pages_per_seq;
  a.page_size = e->page_size;
  a.kv_fp32 = e->kv_fp32;
  a.q_heads = d.q_heads;
  a.kv_heads = d.kv_heads;
  a.head_dim = d.head_dim;
  return a;
}

int batch_tile(int B) { return B <= 1 ? 1 : (B <= 2 ? 2 : (B <= 4 ? 4 : 8)); }

// LM head + greedy argmax over `B` rows of `xin`; finalises next_token / seq_len (+seq_inc)
int lm_head(llmc_engine* e, const float* xin, int64_t x_stride, int B, const int* d_slots, float* logits, int seq_inc,
            int* out_tokens, int* step_counter, bool pdl) {
  GemvArgs a = base_args(e, B);
  const ModelDims& d = e->d;
  a.W = e->lmhead; a.N = d.vocab; a.K = d.hidden;
  a.x = xin; a.x_stride = x_stride; a.gain = e->gf;
  a.slots = d_slots;
  a.y = logits; a.y_stride = d.vocab;
  a.part_val = e->part_val; a.part_idx = e->part_idx; a.counter = e->lm_counter;
  a.next_token_out = e->d_next_token; a.seq_len_out = e->d_seq_len; a.seq_inc = seq_inc;
  a.out_tokens = out_tokens; a.out_stride = MAX_STEPS_PER_CALL; a.step_counter = step_counter;
  E_TRY(launch_gemv_any(EPI_LMHEAD, PRO_NORM, a, pdl, e->stream));
  e->stats.kernel_launches++;
  return LLMC_OK;
}

// One decode step of 9..128 sequences: the matrix products are skinny GEMMs on the tcgen05 kernel of the prefill
// (batch = M tile, every weight streamed once), around them the per-sequence bookkeeping of decode_wide.cu.  Uses
// the prefill workspace (prefill and decode never overlap: both hold the engine mutex).
int enqueue_wide_step(llmc_engine* e, int B, int nsplit) {
  const ModelDims& d = e->d;
  const int H = d.hidden, QD = d.qd(), KD = d.kd(), I = d.inter, V = d.vocab;
  cudaStream_t s = e->stream;
  const bool pdl = e->use_pdl;
  E_TRY(launch_embed_slots(e->emb, e->d_next_token, e->d_slots, B, H, e->px, s));
  for (int l = 0; l < d.layers; ++l) {
    const LayerW& w = e->L[l];
    unsigned char* kp = e->kpool + (size_t)l * e->kv_layer_stride * e->kv_elem;
    unsigned char* vp = e->vpool + (size_t)l * e->kv_layer_stride * e->kv_elem;
    // the norm kernel also clears the QKV buffer, so the split-K GEMM can accumulate into it (no memset node, which
    // would cut the chain of programmatic dependent launches)
    E_TRY(launch_rmsnorm_rows(e->px, w.g1, d.rms_eps, B, H, e->pxn, e->pxn_lo, s, pdl, e->pqkv, QD + 2 * KD));
    if (int rc = gemm(e, e->pxn, e->pxn_lo, w.wqkv, e->pqkv, nullptr, B, QD + 2 * KD, H, GEMM_RESID_F32, true)) return rc;
    {
      AttnDecodeArgs a;
      memset(&a, 0, sizeof(a));
      a.q = e->q; a.qkv_raw = e->pqkv; a.rope_cos = e->rope_cos; a.rope_sin = e->rope_sin; a.out = e->attn; a.out_hi = e->pattn; a.out_lo = e->pattn_lo; a.kpool = kp; a.vpool = vp; a.kv_fp32 = e->kv_fp32;
      a.slots = e->d_slots; a.seq_len = e->d_seq_len; a.ctx_extra = 1;
      a.block_table = e->d_block_table; a.bt_stride = e->pages_per_seq;
      a.page_size = e->page_size; a.q_heads = d.q_heads; a.kv_heads = d.kv_heads; a.head_dim = d.head_dim;
      a.B = B; a.nsplit = nsplit; a.scratch = e->attn_scratch; a.counters = e->attn_counters;
      E_TRY(launch_attn_decode(a, pdl, s));
    }
    if (int rc = gemm(e, e->pattn, e->pattn_lo, w.wo, e->px, nullptr, B, H, QD, GEMM_RESID_F32, true)) return rc;
    E_TRY(launch_rmsnorm_rows(e->px, w.g2, d.rms_eps, B, H, e->pxn, e->pxn_lo, s, pdl));
    if (int rc = gemm(e, e->pxn, e->pxn_lo, w.wgu, e->pact, e->pact_lo, B, 2 * I, H, GEMM_SWIGLU_BF16, true)) return rc;
    if (int rc = gemm(e, e->pact, e->pact_lo, w.wd, e->px, nullptr, B, H, I, GEMM_RESID_F32, true)) return rc;
  }
  E_TRY(launch_rmsnorm_rows(e->px, e->gf, d.rms_eps, B, H, e->pxn, e->pxn_lo, s, pdl));
  if (int rc = gemm(e, e->pxn, e->pxn_lo, e->lmhead, e->logits, nullptr, B, V, H, GEMM_STORE_F32, true)) return rc;
  E_TRY(launch_argmax_finalise(e->logits, B, V, e->d_slots, e->d_next_token, e->d_seq_len, e->d_out_tokens,
                               MAX_STEPS_PER_CALL, e->d_step_counter, e->lm_counter, e->part_val, e->part_idx, s));
  e->stats.kernel_launches += (uint64_t)d.layers * 3 + 4;   // the GEMMs count themselves
  return LLMC_OK;
}

// enqueue one decode step for the B sequences listed in d_slots
int enqueue_decode_step(llmc_engine* e, int B, bool want_logits, int nsplit) {
  if (B >= e->wide_min_batch) return enqueue_wide_step(e, B, nsplit);
  const ModelDims& d = e->d;
  const int H = d.hidden, QD = d.qd(), I = d.inter;
  const bool pdl = e->use_pdl;
  for (int l = 0; l < d.layers; ++l) {
    const LayerW& w = e->L[l];
    unsigned char* kp = e->kpool + (size_t)l * e->kv_layer_stride * e->kv_elem;
    unsigned char* vp = e->vpool + (size_t)l * e->kv_layer_stride * e->kv_elem;
    {  // RMSNorm -> QKV -> RoPE -> KV append   (layer 0 also gathers the embedding)
      GemvArgs a = base_args(e, B);
      a.W = w.wqkv; a.N = d.qkv_rows(); a.K = H;
      a.x = e->x; a.x_stride = H; a.gain = w.g1;
      a.emb = e->emb; a.x_out = e->x;
      a.y = e->q; a.y_stride = QD;
      a.kpool = kp; a.vpool = vp;
      E_TRY(launch_gemv_any(EPI_QKV_ROPE, l == 0 ? PRO_EMBED_NORM : PRO_NORM, a, pdl, e->stream));
    }
    {
      AttnDecodeArgs a;
      memset(&a, 0, sizeof(a));
      a.q = e->q; a.out = e->attn; a.kpool = kp; a.vpool = vp; a.kv_fp32 = e->kv_fp32;
      a.slots = e->d_slots; a.seq_len = e->d_seq_len; a.ctx_extra = 1;
      a.block_table = e->d_block_table; a.bt_stride = e->pages_per_seq;
      a.page_size = e->page_size; a.q_heads = d.q_heads; a.kv_heads = d.kv_heads; a.head_dim = d.head_dim;
      a.B = B; a.nsplit = nsplit; a.scratch = e->attn_scratch; a.counters = e->attn_counters;
      E_TRY(launch_attn_decode(a, pdl, e->stream));
    }
    {  // x += Wo attn
      GemvArgs a = base_args(e, B);
      a.W = w.wo; a.N = H; a.K = QD;
      a.x = e->attn; a.x_stride = QD;
      a.y = e->x; a.y_stride = H;
      E_TRY(launch_gemv_any(EPI_RESID, PRO_PLAIN, a, pdl, e->stream));
    }
    {  // act = silu(Wg h) * (Wu h), h = RMSNorm(x)
      GemvArgs a = base_args(e, B);
      a.W = w.wgu; a.N = 2 * I; a.K = H;
      a.x = e->x; a.x_stride = H; a.gain = w.g2;
      a.y = e->act; a.y_stride = I;
      E_TRY(launch_gemv_any(EPI_SWIGLU, PRO_NORM, a, pdl, e->stream));
    }
    {  // x += Wd act
      GemvArgs a = base_args(e, B);
      a.W = w.wd; a.N = H; a.K = I;
      a.x = e->act; a.x_stride = I;
      a.y = e->x; a.y_stride = H;
      E_TRY(launch_gemv_any(EPI_RESID, PRO_PLAIN, a, pdl, e->stream));
    }
  }
  e->stats.kernel_launches += (uint64_t)d.layers * 5;
  return lm_head(e, e->x, H, B, e->d_slots, want_logits ? e->logits : nullptr, 1, e->d_out_tokens, e->d_step_counter,
                 pdl);
}

int get_graph(llmc_engine* e, int B, bool want_logits, int nsplit, cudaGraphExec_t* out) {
  const int key = (B * 2 + (want_logits ? 1 : 0)) * 16 + nsplit;
  auto it = e->graphs.find(key);
  if (it != e->graphs.end()) {
    *out = it->second;
    return LLMC_OK;
  }
  const uint64_t launches_before = e->stats.kernel_launches;
  E_TRY(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeRelaxed));
  int rc = enqueue_decode_step(e, B, want_logits, nsplit);
  cudaGraph_t g = nullptr;
  cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
  e->stats.kernel_launches = launches_before;  // capture launched nothing
  if (rc != LLMC_OK) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  E_TRY(ce);
  cudaGraphExec_t ge = nullptr;
  ce = cudaGraphInstantiate(&ge, g, 0);
  cudaGraphDestroy(g);
  E_TRY(ce);
  e->graphs[key] = ge;
  *out = ge;
  return LLMC_OK;
}

}  // namespace

// ======================================================================== C ABI
extern "C" {

const char* llmc_version(void) { return "llmc 0.1 (sm_100a)"; }

int llmc_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

const char* llmc_last_error(llmc_engine* e) { return e ? e->last_error.c_str() : g_create_error.c_str(); }

void llmc_engine_destroy(llmc_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (auto& kv : e->graphs) cudaGraphExecDestroy(kv.second);
  for (void* p : e->allocs) cudaFree(p);
  if (e->ev0) cudaEventDestroy(e->ev0);
  if (e->ev1) cudaEventDestroy(e->ev1);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

static int create_impl(llmc_engine* e, const llmc_config& c) {
  const bool init_trace = getenv("LLMC_INIT_TRACE") != nullptr;   // stderr timestamps of the set-up stages
  const auto init_t0 = std::chrono::steady_clock::now();
  auto stage = [&](const char* what) {
    if (init_trace)
      fprintf(stderr, "[init dev %d] %-18s +%.1f ms\n", c.device, what,
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - init_t0).count());
  };
  E_TRY(cudaSetDevice(c.device));
  E_TRY(cudaFree(0));
  stage("context");
  E_TRY(cudaGetDeviceProperties(&e->prop, c.device));
  if (e->prop.major != 10)
    return fail(e, LLMC_ERR_CUDA, std::string("device ") + e->prop.name + " is compute capability " +
                                      std::to_string(e->prop.major) + "." + std::to_string(e->prop.minor) +
                                      "; this library contains sm_100a code only and has no fallback");
  e->device = c.device;
  e->d = ModelDims{c.hidden, c.layers, c.q_heads, c.kv_heads, c.head_dim, c.inter, c.vocab, c.tied,
                   c.rope_theta, c.rms_eps, c.init_std};
  e->seed = c.seed;
  if (c.rope_scaling == LLMC_ROPE_LLAMA3) {
    if (!(c.rope_factor > 0.f) || !(c.rope_low_freq_factor > 0.f) || !(c.rope_high_freq_factor > c.rope_low_freq_factor) ||
        c.rope_original_ctx < 1)
      return fail(e, LLMC_ERR_INVALID, "rope_scaling llama3 needs factor > 0, 0 < low_freq_factor < high_freq_factor and original_ctx >= 1");
    e->rope = RopeScaling{LLMC_ROPE_LLAMA3, c.rope_factor, c.rope_low_freq_factor, c.rope_high_freq_factor, c.rope_original_ctx};
  } else if (c.rope_scaling != 0) {
    return fail(e, LLMC_ERR_INVALID, "unknown rope_scaling id");
  } else if (c.weights_path) {
    std::string err;   // the checkpoint's own config.json decides, when it exists
    if (int rc = read_checkpoint_config(c.weights_path, &e->rope, &e->d.rope_theta, &err)) return fail(e, rc, "weights_path: " + err);
  }
  e->max_slots = c.max_slots; e->max_ctx = c.max_ctx; e->page_size = c.page_size;
  e->prefill_chunk = c.prefill_chunk; e->prefill_impl = c.prefill_impl;
  e->use_graph = !(c.flags & LLMC_FLAG_NO_GRAPH);
  e->use_pdl = !(c.flags & LLMC_FLAG_NO_PDL);
  e->kv_fp32 = (c.flags & LLMC_FLAG_KV_BF16) ? 0 : 1;
  e->kv_elem = e->kv_fp32 ? 4 : 2;
  e->split_prefill = !(c.flags & LLMC_FLAG_PREFILL_BF16);
  e->use_mega = !(c.flags & LLMC_FLAG_NO_MEGAKERNEL);
  {
    // experiment knob: which prefill GEMMs get the hi+lo activation pair (1 QKV, 2 O, 4 gate/up, 8 down); default all
    const char* sm = getenv("LLMC_PREFILL_SPLIT_MASK");
    e->split_mask = sm ? atoi(sm) : 15;
  }
  {
    const char* wm = getenv("LLMC_WIDE_MIN_BATCH");   // experiment knob: smallest batch served by the wide (GEMM) decode path
    e->wide_min_batch = wm ? std::max(2, atoi(wm)) : 2;
    if (e->wide_min_batch > MAX_GEMV_BATCH + 1) e->wide_min_batch = MAX_GEMV_BATCH + 1;
  }
  e->use_ring = e->use_mega && !(c.flags & LLMC_FLAG_NO_RING) &&
                decode_ring_supported(e->d.hidden, e->d.qd(), e->d.kd(), e->d.inter, e->d.vocab, e->d.q_heads, e->d.kv_heads,
                                      e->d.head_dim);
  {
    // LLMC_FLAG_RING_FLOW (or LLMC_RING_FLOW=1) selects the barrier-free variant of the ring kernel (measured slower so
    // far; kept as a comparator)
    const char* fl = getenv("LLMC_RING_FLOW");
    const bool want_flow = fl ? atoi(fl) != 0 : (c.flags & LLMC_FLAG_RING_FLOW) != 0;
    e->use_flow = e->use_ring && want_flow &&
                  decode_flow_supported(e->d.hidden, e->d.qd(), e->d.kd(), e->d.inter, e->d.vocab, e->d.q_heads,
                                        e->d.kv_heads, e->d.head_dim, e->prop.multiProcessorCount);
  }
  {
    const char* rm = getenv("LLMC_RING_MMA");
    const bool want = rm ? atoi(rm) != 0 : (c.flags & LLMC_FLAG_RING_MMA) != 0;
    e->use_ring_mma = e->use_ring && !e->use_flow && want && e->d.hidden / 8 <= 2 * e->prop.multiProcessorCount &&
                      decode_ring_mma_supported(e->d.hidden, e->d.qd(), e->d.kd(), e->d.inter, e->d.vocab, e->d.q_heads,
                                                e->d.kv_heads, e->d.head_dim);
  }
  {
    // 2..4 sequences per decode call: the batched ring kernel unless LLMC_FLAG_NO_RING_BATCH / LLMC_RING_BATCH=0
    const char* rb = getenv("LLMC_RING_BATCH");
    const bool want = rb ? atoi(rb) != 0 : (c.flags & LLMC_FLAG_NO_RING_BATCH) == 0;
    e->use_ringb = e->use_mega && !(c.flags & LLMC_FLAG_NO_RING) && want &&
                   decode_ring_batch_supported(2, e->d.hidden, e->d.qd(), e->d.kd(), e->d.inter, e->d.vocab, e->d.q_heads,
                                               e->d.kv_heads, e->d.head_dim, e->prop.multiProcessorCount);
  }
  e->stats.sm_count = e->prop.multiProcessorCount; e->stats.cc_major = e->prop.major; e->stats.cc_minor = e->prop.minor;
  E_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  E_TRY(cudaEventCreate(&e->ev0));
  E_TRY(cudaEventCreate(&e->ev1));
  if (int rc = init_weights(e, c.weights_path)) return rc;
  if (init_trace) cudaStreamSynchronize(e->stream);
  stage("weights");

  const ModelDims& d = e->d;
  const int H = d.hidden, QD = d.qd(), KD = d.kd(), I = d.inter, V = d.vocab, D = d.head_dim;
  const int MB = MAX_GEMV_BATCH;
  // ---- KV pool: enough pages for every slot at max_ctx
  e->pages_per_seq = (e->max_ctx + e->page_size - 1) / e->page_size;
  e->num_pages = e->pages_per_seq * e->max_slots;
  e->kv_layer_stride = (size_t)e->num_pages * KD * e->page_size;
  if (int rc = dalloc(e, &e->kpool, e->kv_layer_stride * d.layers * e->kv_elem)) return rc;
  if (int rc = dalloc(e, &e->vpool, e->kv_layer_stride * d.layers * e->kv_elem)) return rc;
  if (int rc = dalloc(e, &e->d_block_table, (size_t)e->max_slots * e->pages_per_seq)) return rc;
  E_TRY(cudaMemsetAsync(e->d_block_table, 0, (size_t)e->max_slots * e->pages_per_seq * sizeof(int), e->stream));
  e->h_block_table.assign((size_t)e->max_slots * e->pages_per_seq, 0);
  for (int p = e->num_pages - 1; p >= 0; --p) e->free_pages.push_back(p);
  e->slot_npages.assign(e->max_slots, 0);
  e->page_ref.assign(e->num_pages, 0);
  e->slot_used.assign(e->max_slots, 0);
  e->h_seq_len.assign(e->max_slots, 0);
  if (int rc = dalloc(e, &e->d_seq_len, e->max_slots)) return rc;
  if (int rc = dalloc(e, &e->d_next_token, e->max_slots)) return rc;
  E_TRY(cudaMemsetAsync(e->d_seq_len, 0, e->max_slots * sizeof(int), e->stream));
  E_TRY(cudaMemsetAsync(e->d_next_token, 0, e->max_slots * sizeof(int), e->stream));
  stage("kv pool");
  // ---- RoPE tables (double precision on the host, rounded once to fp32)
  {
    std::vector<float> hc((size_t)e->max_ctx * (D / 2)), hs((size_t)e->max_ctx * (D / 2));
    for (int p = 0; p < e->max_ctx; ++p)
      for (int i = 0; i < D / 2; ++i) {
        double inv = pow((double)d.rope_theta, -2.0 * i / (double)D);
        if (e->rope.kind == LLMC_ROPE_LLAMA3) {
          // transformers' _compute_llama3_parameters: wavelengths beyond original_ctx / low_freq_factor are stretched
          // by `factor`, those below original_ctx / high_freq_factor kept, the band between interpolated
          const double wavelen = 6.283185307179586476925286766559 / inv, oc = (double)e->rope.original_ctx;
          const double lo = (double)e->rope.low_freq_factor, hi = (double)e->rope.high_freq_factor;
          if (wavelen > oc / lo) inv = inv / (double)e->rope.factor;
          else if (!(wavelen < oc / hi)) {
            const double smooth = (oc / wavelen - lo) / (hi - lo);
            inv = (1.0 - smooth) * inv / (double)e->rope.factor + smooth * inv;
          }
        }
        const double a = (double)p * inv;
        hc[(size_t)p * (D / 2) + i] = (float)cos(a);
        hs[(size_t)p * (D / 2) + i] = (float)sin(a);
      }
    if (int rc = dalloc(e, &e->rope_cos, hc.size())) return rc;
    if (int rc = dalloc(e, &e->rope_sin, hs.size())) return rc;
    E_TRY(cudaMemcpy(e->rope_cos, hc.data(), hc.size() * 4, cudaMemcpyHostToDevice));
    E_TRY(cudaMemcpy(e->rope_sin, hs.data(), hs.size() * 4, cudaMemcpyHostToDevice));
  }
  // ---- decode workspace
  // WB: rows of the wide (GEMM) decode path, used when a call carries more than MAX_GEMV_BATCH sequences
  const int WB = std::max(MB, std::min(e->max_slots, MAX_WIDE_BATCH));
  if (int rc = dalloc(e, &e->x, (size_t)MB * H)) return rc;
  if (int rc = dalloc(e, &e->q, (size_t)WB * QD)) return rc;
  if (int rc = dalloc(e, &e->attn, (size_t)WB * QD)) return rc;
  if (int rc = dalloc(e, &e->act, (size_t)MB * I)) return rc;
  if (int rc = dalloc(e, &e->logits, (size_t)WB * V)) return rc;
  if (int rc = dalloc(e, &e->d_slots, WB)) return rc;
  if (int rc = dalloc(e, &e->d_pslot, 1)) return rc;
  const int lmb = std::max(gemv_lmhead_blocks(V), gemv_mma_lmhead_blocks(V));
  const size_t npart = std::max((size_t)MB * lmb, argmax_scratch_elems(WB));
  if (int rc = dalloc(e, &e->part_val, npart)) return rc;
  if (int rc = dalloc(e, &e->part_idx, npart)) return rc;
  if (int rc = dalloc(e, &e->lm_counter, 1)) return rc;
  if (int rc = dalloc(e, &e->attn_counters, (size_t)WB * d.kv_heads)) return rc;
  if (int rc = dalloc(e, &e->attn_scratch, attn_decode_scratch_floats(WB, d.q_heads, ATTN_MAX_SPLIT, D))) return rc;
  if (int rc = dalloc(e, &e->d_out_tokens, (size_t)WB * MAX_STEPS_PER_CALL)) return rc;
  if (int rc = dalloc(e, &e->d_step_counter, 1)) return rc;
  if (int rc = dalloc(e, &e->d_dummy_counter, 1)) return rc;
  E_TRY(cudaMemsetAsync(e->lm_counter, 0, sizeof(unsigned), e->stream));
  E_TRY(cudaMemsetAsync(e->attn_counters, 0, (size_t)WB * d.kv_heads * sizeof(unsigned), e->stream));
  E_TRY(cudaMemsetAsync(e->d_step_counter, 0, sizeof(int), e->stream));
  E_TRY(cudaMemsetAsync(e->d_dummy_counter, 0, sizeof(int), e->stream));
  // ---- one-kernel batch-1 decode
  {
    std::vector<MegaLayer> ml(d.layers);
    for (int l = 0; l < d.layers; ++l) {
      const LayerW& w = e->L[l];
      ml[l] = MegaLayer{w.wqkv, w.wo, w.wgu, w.wd, w.g1, w.g2,
                        e->kpool + (size_t)l * e->kv_layer_stride * e->kv_elem,
                        e->vpool + (size_t)l * e->kv_layer_stride * e->kv_elem};
    }
    if (int rc = dalloc(e, &e->d_mega_layers, (size_t)d.layers)) return rc;
    E_TRY(cudaMemcpy(e->d_mega_layers, ml.data(), ml.size() * sizeof(MegaLayer), cudaMemcpyHostToDevice));
    e->mega_max_tiles = std::max(8, (e->max_ctx + 31) / 32);
    if (int rc = dalloc(e, &e->mega_scratch, (size_t)d.q_heads * e->mega_max_tiles * (D + 2))) return rc;
    if (int rc = dalloc(e, &e->mega_part_val, (size_t)e->prop.multiProcessorCount)) return rc;
    if (int rc = dalloc(e, &e->mega_part_idx, (size_t)e->prop.multiProcessorCount)) return rc;
    if (int rc = dalloc(e, &e->mega_barrier, 4)) return rc;   // [0] grid barrier / pacing counter, [1] attention-output counter
    if (int rc = dalloc(e, &e->mega_counters, (size_t)d.kv_heads)) return rc;
    if (getenv("LLMC_MEGA_TRACE")) {
      const size_t nt = (size_t)2 * (5 * d.layers + 1) * 64 + 16 + 5 * 256;
      if (int rc = dalloc(e, &e->mega_trace, nt)) return rc;
      E_TRY(cudaMemsetAsync(e->mega_trace, 0, nt * sizeof(unsigned long long), e->stream));
    }
    E_TRY(cudaMemsetAsync(e->mega_counters, 0, (size_t)d.kv_heads * sizeof(unsigned), e->stream));
    if (e->use_ringb) {
      const size_t G4 = (size_t)4 * e->prop.multiProcessorCount;
      if (int rc = dalloc(e, &e->ringb_scratch, decode_ring_batch_scratch_floats(d.q_heads, D))) return rc;
      if (int rc = dalloc(e, &e->ringb_part_val, G4)) return rc;
      if (int rc = dalloc(e, &e->ringb_part_idx, G4)) return rc;
      if (int rc = dalloc(e, &e->ringb_counters, (size_t)4 * d.kv_heads)) return rc;
      if (int rc = dalloc(e, &e->ringb_attn_hi, (size_t)4 * QD)) return rc;
      if (int rc = dalloc(e, &e->ringb_attn_lo, (size_t)4 * QD)) return rc;
      if (int rc = dalloc(e, &e->ringb_act_hi, (size_t)4 * I)) return rc;
      if (int rc = dalloc(e, &e->ringb_act_lo, (size_t)4 * I)) return rc;
      if (int rc = dalloc(e, &e->ringb_xn_hi, (size_t)4 * H)) return rc;
      if (int rc = dalloc(e, &e->ringb_xn_lo, (size_t)4 * H)) return rc;
      if (int rc = dalloc(e, &e->ringb_ss_acc, (size_t)16)) return rc;
      if (int rc = dalloc(e, &e->ringb_attn_lo2, (size_t)QD)) return rc;
      if (int rc = dalloc(e, &e->ringb_act_lo2, (size_t)I)) return rc;
      E_TRY(cudaMemsetAsync(e->ringb_counters, 0, (size_t)4 * d.kv_heads * sizeof(unsigned), e->stream));
    }
    if (e->use_flow) {
      e->flow_nwords = decode_flow_words(H, QD, KD, I, d.q_heads, d.kv_heads, D, e->prop.multiProcessorCount);
      if (int rc = dalloc(e, &e->flow_words, e->flow_nwords)) return rc;
      E_TRY(cudaMemsetAsync(e->flow_words, 0, e->flow_nwords * sizeof(unsigned long long), e->stream));
    }
  }
  // ---- prefill workspace
  const int TC = std::max(e->prefill_chunk, WB);   // the wide decode path borrows these buffers (rows = sequences)
  if (int rc = dalloc(e, &e->p_tokens, TC)) return rc;
  if (int rc = dalloc(e, &e->px, (size_t)TC * H)) return rc;
  if (int rc = dalloc(e, &e->pxn, (size_t)TC * H)) return rc;
  if (int rc = dalloc(e, &e->pqkv, (size_t)TC * (QD + 2 * KD))) return rc;
  if (int rc = dalloc(e, &e->pq, (size_t)TC * QD)) return rc;
  if (int rc = dalloc(e, &e->pattn, (size_t)TC * QD)) return rc;
  if (int rc = dalloc(e, &e->pact, (size_t)TC * I)) return rc;
  if (e->split_prefill) {
    if (int rc = dalloc(e, &e->pxn_lo, (size_t)TC * H)) return rc;
    if (int rc = dalloc(e, &e->pattn_lo, (size_t)TC * QD)) return rc;
    if (int rc = dalloc(e, &e->pact_lo, (size_t)TC * I)) return rc;
  }
  if (e->prefill_impl == 0 && prefill_attn_mma_supported(d.head_dim, d.q_heads / d.kv_heads))
    if (int rc = dalloc(e, &e->pattn_ws, prefill_attn_mma_workspace_bytes(e->max_ctx, d.kv_heads, d.head_dim) / 2)) return rc;
  E_TRY(cudaStreamSynchronize(e->stream));
  E_TRY(cudaGetLastError());
  stage("workspaces, sync");
  if (init_trace)
    fprintf(stderr, "[init dev %d] 1..4 sequences: %s\n", c.device, e->use_ringb ? "decode_ring_batch (persistent; batch 1 unless LLMC_RING_BATCH_B1=0 or an experimental ring flag)" : "kernel graph per step (batch 1: below)");
  if (init_trace)
    fprintf(stderr, "[init dev %d] batch-1 decode path: %s\n", c.device,
            !e->use_mega ? "kernel graph" : e->use_flow ? "decode_flow (ring, tagged-word hand-overs)" : e->use_ring_mma ? "decode_ring_mma (ring, mma.sync slots)"
                                          : e->use_ring ? "decode_ring (ring, grid barriers)" : "decode_mega (register streaming)");
  e->stats.graph_enabled = e->use_graph;
  e->stats.pdl_enabled = e->use_pdl;
  return LLMC_OK;
}

int llmc_engine_create(const llmc_config* cfg, llmc_engine** out) {
  if (!cfg || !out) return fail(nullptr, LLMC_ERR_INVALID, "null argument");
  *out = nullptr;
  llmc_config c = *cfg;
  if (apply_shape(c) != LLMC_OK) return fail(nullptr, LLMC_ERR_INVALID, "unknown shape id");
  if (c.rope_theta == 0.f) c.rope_theta = 500000.f;
  if (c.rms_eps == 0.f) c.rms_eps = 1e-5f;
  if (c.init_std == 0.f) c.init_std = 0.02f;
  if (c.max_slots == 0) c.max_slots = 8;
  if (c.max_ctx == 0) c.max_ctx = 8192;
  if (c.page_size == 0) c.page_size = 64;
  if (c.prefill_chunk == 0) c.prefill_chunk = 2048;
  std::string v = validate(c);
  if (!v.empty()) return fail(nullptr, LLMC_ERR_INVALID, v);
  int ndev = llmc_device_count();
  if (ndev <= 0) return fail(nullptr, LLMC_ERR_CUDA, "no CUDA device available; libllmc has no CPU fallback");
  if (c.device < 0 || c.device >= ndev) return fail(nullptr, LLMC_ERR_INVALID, "device ordinal out of range");
  llmc_engine* e = new llmc_engine();
  int rc = create_impl(e, c);
  if (rc != LLMC_OK) {
    g_create_error = e->last_error;
    llmc_engine_destroy(e);
    return rc;
  }
  *out = e;
  return LLMC_OK;
}

int llmc_seq_alloc(llmc_engine* e, int32_t* slot) {
  if (!e || !slot) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  for (int s = 0; s < e->max_slots; ++s)
    if (!e->slot_used[s]) {
      e->slot_used[s] = 1;
      e->h_seq_len[s] = 0;
      *slot = s;
      return LLMC_OK;
    }
  return fail(e, LLMC_ERR_LIMIT, "all " + std::to_string(e->max_slots) + " sequence slots are in use");
}

static int check_slot(llmc_engine* e, int slot) {
  if (slot < 0 || slot >= e->max_slots || !e->slot_used[slot])
    return fail(e, LLMC_ERR_STATE, "slot " + std::to_string(slot) + " is not allocated");
  return LLMC_OK;
}

int llmc_seq_free(llmc_engine* e, int32_t slot) {
  if (!e) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (int rc = check_slot(e, slot)) return rc;
  E_TRY(cudaSetDevice(e->device));
  E_TRY(cudaStreamSynchronize(e->stream));
  int* row = &e->h_block_table[(size_t)slot * e->pages_per_seq];
  for (int i = 0; i < e->slot_npages[slot]; ++i)
    if (--e->page_ref[row[i]] == 0) e->free_pages.push_back(row[i]);   // pages shared with a fork live on
  e->slot_npages[slot] = 0;
  e->slot_used[slot] = 0;
  e->h_seq_len[slot] = 0;
  int zero = 0;
  E_TRY(cudaMemcpyAsync(e->d_seq_len + slot, &zero, sizeof(int), cudaMemcpyHostToDevice, e->stream));
  E_TRY(cudaStreamSynchronize(e->stream));
  return LLMC_OK;
}

int llmc_seq_fork(llmc_engine* e, int32_t src, int32_t n_tokens, int32_t* new_slot) {
  if (!e || !new_slot) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (int rc = check_slot(e, src)) return rc;
  if (n_tokens < 1 || n_tokens > e->h_seq_len[src])
    return fail(e, LLMC_ERR_INVALID, "llmc_seq_fork: n_tokens must be in [1, " + std::to_string(e->h_seq_len[src]) + "]");
  int dst = -1;
  for (int s = 0; s < e->max_slots; ++s)
    if (!e->slot_used[s]) { dst = s; break; }
  if (dst < 0) return fail(e, LLMC_ERR_LIMIT, "all " + std::to_string(e->max_slots) + " sequence slots are in use");
  const int full = n_tokens / e->page_size, rem = n_tokens % e->page_size;
  if (rem && e->free_pages.empty()) return fail(e, LLMC_ERR_OOM, "KV page pool exhausted");
  E_TRY(cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  const int* srow = &e->h_block_table[(size_t)src * e->pages_per_seq];
  int* drow = &e->h_block_table[(size_t)dst * e->pages_per_seq];
  // full pages are never written again (both sequences append behind them): shared.  The page that holds the last
  // n_tokens % page_size positions is copied, all layers, K and V: one strided copy per pool.
  for (int i = 0; i < full; ++i) {
    drow[i] = srow[i];
    ++e->page_ref[srow[i]];
  }
  int npages = full;
  if (rem) {
    const int pg = e->free_pages.back();
    e->free_pages.pop_back();
    e->page_ref[pg] = 1;
    drow[full] = pg;
    npages = full + 1;
    const size_t page_bytes = (size_t)e->d.kd() * e->page_size * e->kv_elem;
    const size_t pitch = e->kv_layer_stride * e->kv_elem;
    for (unsigned char* pool : {e->kpool, e->vpool}) {
      unsigned char* to = pool + (size_t)pg * page_bytes;
      const unsigned char* from = pool + (size_t)srow[full] * page_bytes;
      if (pitch <= e->prop.memPitch) {
        E_TRY(cudaMemcpy2DAsync(to, pitch, from, pitch, page_bytes, e->d.layers, cudaMemcpyDeviceToDevice, s));
      } else {   // very large pools: the 2-D copy's pitch limit
        for (int l = 0; l < e->d.layers; ++l)
          E_TRY(cudaMemcpyAsync(to + (size_t)l * pitch, from + (size_t)l * pitch, page_bytes, cudaMemcpyDeviceToDevice, s));
      }
    }
  }
  e->slot_used[dst] = 1;
  e->slot_npages[dst] = npages;
  e->h_seq_len[dst] = n_tokens;
  E_TRY(cudaMemcpyAsync(e->d_block_table + (size_t)dst * e->pages_per_seq, drow, (size_t)npages * sizeof(int),
                        cudaMemcpyHostToDevice, s));
  E_TRY(cudaMemcpyAsync(e->d_seq_len + dst, &e->h_seq_len[dst], sizeof(int), cudaMemcpyHostToDevice, s));
  if (n_tokens == e->h_seq_len[src])   // a clone at the source's end also inherits its pending token
    E_TRY(cudaMemcpyAsync(e->d_next_token + dst, e->d_next_token + src, sizeof(int), cudaMemcpyDeviceToDevice, s));
  E_TRY(cudaStreamSynchronize(s));
  *new_slot = dst;
  return LLMC_OK;
}

int llmc_seq_len(llmc_engine* e, int32_t slot, int32_t* len) {
  if (!e || !len) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (int rc = check_slot(e, slot)) return rc;
  *len = e->h_seq_len[slot];
  return LLMC_OK;
}

int llmc_seq_set_next_token(llmc_engine* e, int32_t slot, int32_t token) {
  if (!e) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (int rc = check_slot(e, slot)) return rc;
  if (token < 0 || token >= e->d.vocab) return fail(e, LLMC_ERR_INVALID, "token id out of range");
  E_TRY(cudaSetDevice(e->device));
  E_TRY(cudaMemcpyAsync(e->d_next_token + slot, &token, sizeof(int), cudaMemcpyHostToDevice, e->stream));
  E_TRY(cudaStreamSynchronize(e->stream));
  return LLMC_OK;
}

int llmc_seq_get_next_token(llmc_engine* e, int32_t slot, int32_t* token) {
  if (!e || !token) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (int rc = check_slot(e, slot)) return rc;
  E_TRY(cudaSetDevice(e->device));
  E_TRY(cudaMemcpyAsync(token, e->d_next_token + slot, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
  E_TRY(cudaStreamSynchronize(e->stream));
  return LLMC_OK;
}

int llmc_prefill(llmc_engine* e, int32_t slot, const int32_t* tokens, int32_t n, int32_t* next_token,
                 float* logits_last) {
  if (!e || !tokens || n < 1) return e ? fail(e, LLMC_ERR_INVALID, "llmc_prefill: null tokens or n < 1") : LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (int rc = check_slot(e, slot)) return rc;
  const ModelDims& d = e->d;
  for (int i = 0; i < n; ++i)
    if (tokens[i] < 0 || tokens[i] >= d.vocab) return fail(e, LLMC_ERR_INVALID, "token id out of range");
  E_TRY(cudaSetDevice(e->device));
  const int H = d.hidden, QD = d.qd(), KD = d.kd(), I = d.inter;
  const int p_start = e->h_seq_len[slot];
  if (int rc = ensure_pages(e, slot, p_start + n)) return rc;
  cudaStream_t s = e->stream;
  E_TRY(cudaEventRecord(e->ev0, s));
  int last_tc = 0;
  for (int c0 = 0; c0 < n; c0 += e->prefill_chunk) {
    const int T = std::min(e->prefill_chunk, n - c0);
    const int p0 = p_start + c0;
    last_tc = T;
    E_TRY(cudaMemcpyAsync(e->p_tokens, tokens + c0, (size_t)T * sizeof(int), cudaMemcpyHostToDevice, s));
    E_TRY(launch_embed_rows(e->emb, e->p_tokens, T, H, e->px, s));
    for (int l = 0; l < d.layers; ++l) {
      const LayerW& w = e->L[l];
      unsigned char* kp = e->kpool + (size_t)l * e->kv_layer_stride * e->kv_elem;
      unsigned char* vp = e->vpool + (size_t)l * e->kv_layer_stride * e->kv_elem;
      E_TRY(launch_rmsnorm_rows(e->px, w.g1, d.rms_eps, T, H, e->pxn, e->pxn_lo, s));
      if (int rc = gemm(e, e->pxn, (e->split_mask & 1) ? e->pxn_lo : nullptr, w.wqkv, e->pqkv, nullptr, T, QD + 2 * KD, H, GEMM_STORE_F32)) return rc;
      E_TRY(launch_rope_append(e->pqkv, T, p0, slot, e->rope_cos, e->rope_sin, e->d_block_table, e->pages_per_seq,
                               kp, vp, e->kv_fp32, e->page_size, d.q_heads, d.kv_heads, d.head_dim, e->pq, s));
      E_TRY(launch_prefill_attention(e->pq, T, p0, slot, e->d_block_table, e->pages_per_seq, kp, vp, e->kv_fp32,
                                     e->page_size, d.q_heads, d.kv_heads, d.head_dim, e->pattn, e->pattn_lo, e->pattn_ws,
                                     e->prefill_impl, s));
      if (int rc = gemm(e, e->pattn, (e->split_mask & 2) ? e->pattn_lo : nullptr, w.wo, e->px, nullptr, T, H, QD, GEMM_RESID_F32)) return rc;
      E_TRY(launch_rmsnorm_rows(e->px, w.g2, d.rms_eps, T, H, e->pxn, e->pxn_lo, s));
      if (int rc = gemm(e, e->pxn, (e->split_mask & 4) ? e->pxn_lo : nullptr, w.wgu, e->pact, e->pact_lo, T, 2 * I, H, GEMM_SWIGLU_BF16)) return rc;
      if (int rc = gemm(e, e->pact, (e->split_mask & 8) ? e->pact_lo : nullptr, w.wd, e->px, nullptr, T, H, I, GEMM_RESID_F32)) return rc;
    }
    e->stats.kernel_launches += 1 + (uint64_t)d.layers * 4;
  }
  // the cache now holds p_start + n tokens; LM head + argmax on the last row
  const int new_len = p_start + n;
  E_TRY(cudaMemcpyAsync(e->d_seq_len + slot, &new_len, sizeof(int), cudaMemcpyHostToDevice, s));
  int hslot = slot;
  E_TRY(cudaMemcpyAsync(e->d_pslot, &hslot, sizeof(int), cudaMemcpyHostToDevice, s));
  if (int rc = lm_head(e, e->px + (size_t)(last_tc - 1) * H, H, 1, e->d_pslot, logits_last ? e->logits : nullptr, 0,
                       nullptr, e->d_dummy_counter, false))
    return rc;
  E_TRY(cudaEventRecord(e->ev1, s));
  int tok = -1;
  E_TRY(cudaMemcpyAsync(&tok, e->d_next_token + slot, sizeof(int), cudaMemcpyDeviceToHost, s));
  if (logits_last)
    E_TRY(cudaMemcpyAsync(logits_last, e->logits, (size_t)d.vocab * sizeof(float), cudaMemcpyDeviceToHost, s));
  E_TRY(cudaStreamSynchronize(s));
  E_TRY(cudaGetLastError());
  e->h_seq_len[slot] = new_len;
  if (next_token) *next_token = tok;
  e->stats.prefill_tokens += n;
  cudaEventElapsedTime(&e->stats.last_prefill_ms, e->ev0, e->ev1);
  return LLMC_OK;
}

// one group of <= MAX_GEMV_BATCH sequences decoded together (engine mutex held by the caller)
static int decode_group(llmc_engine* e, const int32_t* slots, int32_t nslots, int32_t nsteps, int32_t* out_tokens,
                        float* logits) {
  E_TRY(cudaSetDevice(e->device));
  for (int i = 0; i < nslots; ++i)
    if (int rc = ensure_pages(e, slots[i], e->h_seq_len[slots[i]] + nsteps)) return rc;
  cudaStream_t s = e->stream;
  const ModelDims& d = e->d;
  E_TRY(cudaMemcpyAsync(e->d_slots, slots, (size_t)nslots * sizeof(int), cudaMemcpyHostToDevice, s));
  E_TRY(cudaMemsetAsync(e->d_step_counter, 0, sizeof(int), s));
  const bool want_logits = logits != nullptr;
  int max_len = 0;
  for (int i = 0; i < nslots; ++i) max_len = std::max(max_len, e->h_seq_len[slots[i]]);
  // blocks per (sequence, kv head) in the attention kernel, fixed for this call from the longest
  // context it will reach (the graph is keyed by it)
  const int nsplit = attn_decode_nsplit(max_len + nsteps);
  cudaGraphExec_t ge = nullptr;
  // batch 1 takes the batched kernel too (second sequence absent): 0.632 ms per step at the 1B shape against 0.655 for
  // decode_ring.cu; LLMC_RING_BATCH_B1=0 (or the experimental ring flags) keeps batch 1 on the single-sequence kernels
  static const bool ringb_b1 = !(getenv("LLMC_RING_BATCH_B1") && atoi(getenv("LLMC_RING_BATCH_B1")) == 0);
  const bool b1_ok = ringb_b1 && !e->use_flow && !e->use_ring_mma;
  const bool ringb = e->use_ringb && nslots >= (b1_ok ? 1 : 2) && nslots <= 4 && decode_ring_batch_supported(nslots, d.hidden, d.qd(), d.kd(), d.inter, d.vocab, d.q_heads, d.kv_heads,
                                                 d.head_dim, e->prop.multiProcessorCount);
  if (e->use_graph && !(e->use_mega && nslots == 1) && !ringb) {
    int rc = get_graph(e, nslots, want_logits, nsplit, &ge);
    if (rc != LLMC_OK) return rc;
  }
  const bool mega = e->use_mega && nslots == 1 && !ringb;
  if (mega) ge = nullptr;
  E_TRY(cudaEventRecord(e->ev0, s));
  if (ringb) {
    MegaArgs m;
    memset(&m, 0, sizeof(m));
    m.layers = e->d_mega_layers; m.L = d.layers; m.H = d.hidden; m.QD = d.qd(); m.KD = d.kd(); m.I = d.inter;
    m.V = d.vocab; m.q_heads = d.q_heads; m.kv_heads = d.kv_heads; m.head_dim = d.head_dim; m.eps = d.rms_eps;
    m.emb = e->emb; m.lmhead = e->lmhead; m.gf = e->gf; m.rope_cos = e->rope_cos; m.rope_sin = e->rope_sin;
    m.block_table = e->d_block_table; m.bt_stride = e->pages_per_seq; m.page_size = e->page_size;
    m.kv_fp32 = e->kv_fp32; m.seq_len = e->d_seq_len; m.next_token = e->d_next_token;
    m.x = e->x; m.q = e->q; m.attn = e->attn; m.act = e->act; m.logits = want_logits ? e->logits : nullptr;
    m.attn_scratch = e->ringb_scratch; m.max_tiles = e->mega_max_tiles; m.attn_counters = e->ringb_counters;
    m.part_val = e->ringb_part_val; m.part_idx = e->ringb_part_idx; m.out_tokens = e->d_out_tokens; m.nsteps = nsteps;
    m.out_stride = MAX_STEPS_PER_CALL; m.slots = e->d_slots; m.B = nslots;
    m.attn_hi = e->ringb_attn_hi; m.attn_lo = e->ringb_attn_lo; m.act_hi = e->ringb_act_hi; m.act_lo = e->ringb_act_lo;
    m.attn_lo2 = e->ringb_attn_lo2; m.act_lo2 = e->ringb_act_lo2;
    m.xn_hi = e->ringb_xn_hi; m.xn_lo = e->ringb_xn_lo; m.ss_acc = e->ringb_ss_acc;
    E_TRY(cudaMemsetAsync(e->ringb_ss_acc, 0, 16 * sizeof(unsigned long long), s));
    m.barrier = e->mega_barrier;
    {
      const char* tpb = getenv("LLMC_MEGA_TILES_PER_BLOCK");
      m.attn_tiles_per_block = tpb ? std::max(1, atoi(tpb)) : 4;
    }
    m.trace = (e->mega_trace && nsteps <= 64) ? e->mega_trace : nullptr;
    E_TRY(cudaMemsetAsync(e->mega_barrier, 0, 4 * sizeof(unsigned), s));
    E_TRY(launch_decode_ring_batch(m, e->prop.multiProcessorCount, s));
    e->stats.kernel_launches += 1;
  }
  if (mega) {
    MegaArgs m;
    memset(&m, 0, sizeof(m));
    m.layers = e->d_mega_layers; m.L = d.layers; m.H = d.hidden; m.QD = d.qd(); m.KD = d.kd(); m.I = d.inter;
    m.V = d.vocab; m.q_heads = d.q_heads; m.kv_heads = d.kv_heads; m.head_dim = d.head_dim; m.eps = d.rms_eps;
    m.emb = e->emb; m.lmhead = e->lmhead; m.gf = e->gf; m.rope_cos = e->rope_cos; m.rope_sin = e->rope_sin;
    m.block_table = e->d_block_table; m.slot = slots[0]; m.bt_stride = e->pages_per_seq; m.page_size = e->page_size;
    m.kv_fp32 = e->kv_fp32; m.seq_len = e->d_seq_len; m.next_token = e->d_next_token;
    m.x = e->x; m.q = e->q; m.attn = e->attn; m.act = e->act; m.logits = want_logits ? e->logits : nullptr;
    m.attn_scratch = e->mega_scratch; m.max_tiles = e->mega_max_tiles; m.attn_counters = e->mega_counters;
    m.part_val = e->mega_part_val; m.part_idx = e->mega_part_idx; m.out_tokens = e->d_out_tokens; m.nsteps = nsteps;
    m.barrier = e->mega_barrier;
    {
      const char* tpb = getenv("LLMC_MEGA_TILES_PER_BLOCK");
      m.attn_tiles_per_block = tpb ? std::max(1, atoi(tpb)) : 4;
    }
    {
      const char* pm = getenv("LLMC_RING_POLL");
      m.ring_poll_mode = pm ? atoi(pm) : 1;
    }
    m.trace = (e->mega_trace && nsteps <= 64) ? e->mega_trace : nullptr;
    E_TRY(cudaMemsetAsync(e->mega_barrier, 0, 4 * sizeof(unsigned), s));
    if (e->use_flow) {
      // hand-over tags continue from launch to launch, so a word left by an earlier call never matches
      const unsigned need = (unsigned)nsteps * (unsigned)(5 * d.layers + 1);
      if (e->flow_tag > 0xF0000000u - need) {   // before the 32-bit tags wrap: clear the words, restart at 1
        E_TRY(cudaMemsetAsync(e->flow_words, 0, e->flow_nwords * sizeof(unsigned long long), s));
        e->flow_tag = 1;
      }
      m.tw = e->flow_words;
      m.tag0 = e->flow_tag;
      e->flow_tag += need;
      const char* fp = getenv("LLMC_FLOW_POLL");
      m.flow_poll_mode = fp ? atoi(fp) : 0;
      E_TRY(launch_decode_flow(m, e->prop.multiProcessorCount, s));
    } else if (e->use_ring_mma) E_TRY(launch_decode_ring_mma(m, e->prop.multiProcessorCount, s));
    else if (e->use_ring) E_TRY(launch_decode_ring(m, e->prop.multiProcessorCount, s));
    else E_TRY(launch_decode_mega(m, e->prop.multiProcessorCount, s));
    e->stats.kernel_launches += 1;
  }
  for (int st = 0; st < nsteps && !mega && !ringb; ++st) {
    if (ge) {
      E_TRY(cudaGraphLaunch(ge, s));
      e->stats.kernel_launches += nslots >= e->wide_min_batch ? (uint64_t)d.layers * 7 + 5 : (uint64_t)d.layers * 5 + 1;
    } else {
      if (int rc = enqueue_decode_step(e, nslots, want_logits, nsplit)) return rc;
    }
  }
  E_TRY(cudaEventRecord(e->ev1, s));
  std::vector<int> host((size_t)nslots * nsteps);
  for (int i = 0; i < nslots; ++i)
    E_TRY(cudaMemcpyAsync(host.data() + (size_t)i * nsteps, e->d_out_tokens + (size_t)i * MAX_STEPS_PER_CALL,
                          (size_t)nsteps * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (want_logits)
    E_TRY(cudaMemcpyAsync(logits, e->logits, (size_t)nslots * d.vocab * sizeof(float), cudaMemcpyDeviceToHost, s));
  E_TRY(cudaStreamSynchronize(s));
  E_TRY(cudaGetLastError());
  if ((mega || ringb) && e->mega_trace && nsteps <= 64) {
    // debug trace: average per-phase time (barrier open -> next barrier open) and barrier wait of block 0
    const int nb = (5 * d.layers + 1) * nsteps;
    std::vector<unsigned long long> t((size_t)2 * nb + 16 + 5 * 256);
    E_TRY(cudaMemcpy(t.data(), e->mega_trace, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    const char* names[6] = {"qkv", "attn", "o", "gateup", "down", "lmhead"};
    double phase[6] = {0, 0, 0, 0, 0, 0}, wait[6] = {0, 0, 0, 0, 0, 0};
    int cnt[6] = {0, 0, 0, 0, 0, 0};
    const int per = 5 * d.layers + 1;
    for (int i = 1; i < nb; ++i) {
      const int k = (i % per) == per - 1 ? 5 : (i % per) % 5;
      phase[k] += (double)(t[2 * i + 1] - t[2 * (i - 1) + 1]);
      wait[k] += (double)(t[2 * i + 1] - t[2 * i]);
      cnt[k]++;
    }
    for (int k = 0; k < 6; ++k)
      if (cnt[k]) fprintf(stderr, "[%s trace] %-7s phase %.2f us (block 0 waited %.2f us at its barrier) x%d\n", ringb ? "ring batch" : e->use_flow ? "flow" : e->use_ring ? "ring" : "mega", names[k],
                          phase[k] / cnt[k] / 1e3, wait[k] / cnt[k] / 1e3, cnt[k]);
    if (ringb) {
      const unsigned long long* st = t.data() + (size_t)2 * nb;
      fprintf(stderr, "[ring batch stats] block 0 warp 0, cycles: total %llu = wait-full %llu + activation loads %llu + slots %llu "
                      "(%llu slots seen; K = 2048 phases only) + exchanges %llu + grid barriers %llu + rest\n",
              st[4], st[0], st[1], st[2], st[5], st[6], st[3]);
    }
    if (mega && e->use_flow) {
      const unsigned long long* st = t.data() + (size_t)2 * nb;
      fprintf(stderr, "[flow stats] block 0 warp 0, cycles: total %llu = wait-full %llu + activation loads %llu + slots %llu "
                      "(%llu slots seen) + hand-overs (all-writer) %llu + rest\n", st[4], st[0], st[1], st[2], st[5], st[3]);
    }
    if (mega && e->use_ring && !e->use_flow) {
      const unsigned long long* st = t.data() + (size_t)2 * nb;
      if (nsteps > 1) {
        const int G = e->prop.multiProcessorCount;
        const char* pn[5] = {"qkv", "attn", "o", "gateup", "down"};
        for (int k = 0; k < 5; ++k) {
          std::vector<unsigned long long> v(st + 16 + (size_t)k * G, st + 16 + (size_t)(k + 1) * G);
          std::sort(v.begin(), v.end());
          const double last = (double)v[G - 1];
          fprintf(stderr, "[ring arrivals] %-6s barrier of step 1 layer 1: first block %.2f us before the last, 10%% %.2f, median %.2f, "
                          "90%% %.2f, second-to-last %.2f\n", pn[k], (last - v[0]) / 1e3, (last - v[G / 10]) / 1e3,
                  (last - v[G / 2]) / 1e3, (last - v[G * 9 / 10]) / 1e3, (last - v[G - 2]) / 1e3);
        }
      }
      fprintf(stderr, "[ring stats] block 0 warp 0, cycles: total %llu = wait-full %llu + x-load %llu + slots %llu (%llu slots seen) "
                      "+ grid barriers %llu; producer: total %llu, wait-empty %llu, %llu slots\n",
              st[4], st[0], st[1], st[2], st[5], st[3], st[9], st[8], st[10]);
    }
  }
  if (out_tokens) memcpy(out_tokens, host.data(), host.size() * sizeof(int));
  for (int i = 0; i < nslots; ++i) e->h_seq_len[slots[i]] += nsteps;
  cudaEventElapsedTime(&e->stats.last_decode_ms, e->ev0, e->ev1);
  return LLMC_OK;
}

int llmc_decode(llmc_engine* e, const int32_t* slots, int32_t nslots, int32_t nsteps, int32_t* out_tokens,
                float* logits) {
  if (!e || !slots) return e ? fail(e, LLMC_ERR_INVALID, "llmc_decode: null slots") : LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  if (nslots < 1 || nslots > e->max_slots)
    return fail(e, LLMC_ERR_LIMIT, "llmc_decode: nslots must be in [1," + std::to_string(e->max_slots) + "]");
  if (nsteps < 1 || nsteps > MAX_STEPS_PER_CALL)
    return fail(e, LLMC_ERR_LIMIT, "llmc_decode: nsteps must be in [1," + std::to_string(MAX_STEPS_PER_CALL) + "]");
  for (int i = 0; i < nslots; ++i) {
    if (int rc = check_slot(e, slots[i])) return rc;
    if (e->h_seq_len[slots[i]] < 1) return fail(e, LLMC_ERR_STATE, "llmc_decode: slot has no prefilled tokens");
    for (int j = 0; j < i; ++j)
      if (slots[j] == slots[i]) return fail(e, LLMC_ERR_INVALID, "llmc_decode: duplicate slot");
    if (e->h_seq_len[slots[i]] + nsteps > e->max_ctx)
      return fail(e, LLMC_ERR_LIMIT, "sequence would exceed max_ctx (" + std::to_string(e->max_ctx) + " tokens)");
  }
  // Up to 8 sequences: the GEMV kernels; 9..128: the wide path (tcgen05 skinny GEMMs), one pass over the
  // weights for the whole group; more than 128 run as consecutive groups (sequences are independent).
  float total_ms = 0.f;
  const int group = nslots >= e->wide_min_batch ? MAX_WIDE_BATCH : MAX_GEMV_BATCH;
  for (int g0 = 0; g0 < nslots; g0 += group) {
    const int n = std::min<int>(group, nslots - g0);
    int rc = decode_group(e, slots + g0, n, nsteps, out_tokens ? out_tokens + (size_t)g0 * nsteps : nullptr,
                          logits ? logits + (size_t)g0 * e->d.vocab : nullptr);
    if (rc != LLMC_OK) return rc;
    total_ms += e->stats.last_decode_ms;
  }
  e->stats.last_decode_ms = total_ms;
  e->stats.decode_steps += nsteps;
  return LLMC_OK;
}

int llmc_engine_stats(llmc_engine* e, llmc_stats* out) {
  if (!e || !out) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  *out = e->stats;
  return LLMC_OK;
}

}  // extern "C"
