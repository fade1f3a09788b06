/* llmc.h -- C ABI of libllmc.so, the B200-native local inference library that sits behind
 * llm-consensus' provider interface.
 *
 * What this boundary replaces.  In johnayoung/llm-consensus every model is reached through
 *     Provider.QueryStream(ctx, Request{Model, Prompt}, StreamCallback) (Response, error)
 *         /root/reference/internal/provider/provider.go:13-21
 * and the three shipped backends implement it with an HTTPS POST plus an SSE read loop
 * (/root/reference/internal/provider/openai.go:141-210, anthropic.go:132-202,
 * google.go:136-208).  A local backend implements the same two methods by tokenising the
 * prompt and driving THIS library: llmc_prefill() stands where the HTTP request is sent,
 * the llmc_decode() loop stands where the SSE `data:` lines are scanned and handed to the
 * callback.  The reference has no FFI of its own (CGO_ENABLED=0, .goreleaser.yaml:15), so
 * these entry points are what its cgo binding for the path would bind; INTEGRATION.md
 * shows that binding.
 *
 * Conventions:
 *   - plain C types only; no C++ exceptions cross the boundary; nothing calls abort()/exit();
 *   - every function returns 0 on success or a negative LLMC_ERR_* code; the message is
 *     available from llmc_last_error() (per engine; NULL engine = the creating thread's
 *     last create/destroy error);
 *   - the caller owns every host buffer it passes; the library owns all device memory and
 *     frees it in llmc_engine_destroy(); no pointer is retained across calls; no callbacks;
 *   - every entry point selects its engine's device itself, so it may be called from any
 *     OS thread (goroutines migrate).  Calls on ONE engine serialise on an internal mutex;
 *     engines on different GPUs run concurrently with no shared lock
 *     (runner fan-out, /root/reference/internal/runner/runner.go:62-95);
 *   - there is no CPU fallback: without a CUDA device (or on a device that is not
 *     sm_100) llmc_engine_create() fails with LLMC_ERR_CUDA.
 */
#ifndef LLMC_H
#define LLMC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LLMC_OK 0
#define LLMC_ERR_INVALID (-1) /* bad argument / unsupported shape */
#define LLMC_ERR_CUDA (-2)    /* CUDA runtime/driver error, message has the detail */
#define LLMC_ERR_OOM (-3)     /* device memory or KV pages exhausted */
#define LLMC_ERR_STATE (-4)   /* slot not allocated, nothing prefilled, ... */
#define LLMC_ERR_LIMIT (-5)   /* a stated maximum would be exceeded */

/* model shapes named by BASELINE.json */
#define LLMC_SHAPE_CUSTOM 0
#define LLMC_SHAPE_LLAMA_1B 1 /* H 2048, L 16, 32q/8kv x 64, I 8192, V 128256, tied   */
#define LLMC_SHAPE_LLAMA_8B 2 /* H 4096, L 32, 32q/8kv x 128, I 14336, V 128256, untied */

typedef struct llmc_engine llmc_engine;

typedef struct llmc_config {
  int32_t device;       /* CUDA ordinal */
  int32_t shape;        /* LLMC_SHAPE_*; CUSTOM uses the explicit dims below */
  int32_t hidden, layers, q_heads, kv_heads, head_dim, inter, vocab, tied;
  float rope_theta;     /* 0 -> 500000 */
  float rms_eps;        /* 0 -> 1e-5 */
  float init_std;       /* 0 -> 0.02; weights are bf16(std*sqrt(3)*u), u from the counter RNG */
  uint64_t seed;        /* weight seed (the provider's --seed) */
  const char* weights_path; /* NULL: random init from `seed`.  Else a .safetensors file or a directory of shards with
                             * Hugging Face Llama tensor names (BF16 / F16 / F32; rounded to bf16), matching the shape. */
  int32_t max_slots;    /* concurrent sequences (0 -> 8) */
  int32_t max_ctx;      /* tokens per sequence (0 -> 8192) */
  int32_t page_size;    /* KV page, tokens; multiple of 32 (0 -> 64) */
  int32_t prefill_chunk;/* tokens per prefill pass (0 -> 2048) */
  int32_t prefill_impl; /* 0: tcgen05 tensor-core GEMM (product path); 1: plain CUDA-core GEMM (bring-up/comparator) */
  int32_t flags;        /* LLMC_FLAG_* */
  /* RoPE frequency scaling, Hugging Face rope_type "llama3" (what Llama-3.1 / 3.2 checkpoints declare in
   * config.json): 0 = none, LLMC_ROPE_LLAMA3 = rescale inv_freq with the four parameters.  With weights_path set
   * and rope_scaling == 0, a config.json beside the checkpoint is read instead (rope_theta and rope_scaling);
   * a rope_scaling type other than "default" / "llama3" is refused. */
  int32_t rope_scaling;
  float rope_factor, rope_low_freq_factor, rope_high_freq_factor;
  int32_t rope_original_ctx; /* original_max_position_embeddings */
} llmc_config;
#define LLMC_ROPE_LLAMA3 1

#define LLMC_FLAG_NO_GRAPH 1 /* launch decode kernels one by one instead of as a CUDA graph */
#define LLMC_FLAG_NO_PDL 2   /* no programmatic dependent launch between decode kernels */
#define LLMC_FLAG_KV_BF16 4  /* store the KV cache in bf16 (default fp32; see DESIGN.md "precision") */
#define LLMC_FLAG_PREFILL_BF16 8 /* prefill GEMM activations as one bf16 (default: a bf16 hi+lo pair) */
#define LLMC_FLAG_NO_MEGAKERNEL 16 /* batch-1 decode as 81 kernels per step instead of one persistent kernel */
#define LLMC_FLAG_RING_FLOW 64 /* experimental: ring kernel with tagged-word hand-overs (decode_flow.cu) instead of grid barriers between phases (decode_ring.cu) */
#define LLMC_FLAG_RING_MMA 256 /* experimental: ring kernel with the slot arithmetic on mma.sync (decode_ring_mma.cu); same speed as the FMA loop, kept as a comparator */
#define LLMC_FLAG_NO_RING_BATCH 128 /* no decode_ring_batch.cu: batch 1 on decode_ring.cu (or decode_mega.cu), 2..4 sequences through the per-step kernel graph */
#define LLMC_FLAG_NO_RING 32 /* persistent batch-1 decode pulls weights through registers (decode_mega.cu) instead of the TMA shared-memory ring (decode_ring.cu) */

typedef struct llmc_stats {
  uint64_t kernel_launches;  /* kernels of this library launched since create (graph nodes counted) */
  uint64_t decode_steps;
  uint64_t prefill_tokens;
  double weight_bytes;       /* bf16 bytes streamed per decode step: layers + LM head + gains */
  double kv_bytes_per_token; /* 2 * kv_heads * head_dim * layers * (4 B fp32 | 2 B bf16) */
  float last_prefill_ms;     /* CUDA-event time of the last llmc_prefill on the engine's stream */
  float last_decode_ms;      /* CUDA-event time of the last llmc_decode (all its steps) */
  int32_t graph_enabled, pdl_enabled;
  int32_t sm_count, cc_major, cc_minor;
} llmc_stats;

/* Number of CUDA devices visible; <= 0 means the library cannot run here. */
int llmc_device_count(void);

/* Build an engine on cfg->device: allocate weights, generate them from the seed on the
 * device, allocate the paged KV pool and work buffers.  One engine = one model instance
 * (createProvider, /root/reference/cmd/llm-consensus/main.go:417-438, is called once per
 * unique model name). */
int llmc_engine_create(const llmc_config* cfg, llmc_engine** out);
void llmc_engine_destroy(llmc_engine* e);

/* Sequence slots: one per in-flight QueryStream call. */
int llmc_seq_alloc(llmc_engine* e, int32_t* slot);
int llmc_seq_free(llmc_engine* e, int32_t slot);
/* New slot holding the first n_tokens (1..llmc_seq_len(src)) cached tokens of slot src: full KV pages are shared
 * read-only (reference-counted; both sequences only ever append behind them), the page holding the last
 * n_tokens % page_size positions is copied.  n_tokens == llmc_seq_len(src) also copies the pending token, so the fork
 * decodes exactly what src would.  This is how the provider serves what the reference sends it repeatedly: the same
 * prompt once per duplicate --models entry (/root/reference/internal/runner/runner.go:62-95 starts one goroutine per
 * entry, no dedupe) and the judge prompt's fixed head (/root/reference/internal/consensus/judge.go:12-19, 96-99). */
int llmc_seq_fork(llmc_engine* e, int32_t src, int32_t n_tokens, int32_t* new_slot);
/* tokens currently held in the slot's KV cache */
int llmc_seq_len(llmc_engine* e, int32_t slot, int32_t* len);

/* Append n prompt tokens to the slot and run them through the model (chunked).  The greedy
 * token following the prompt becomes the slot's pending token and is returned in
 * *next_token (may be NULL); logits_last (may be NULL) receives the vocab fp32 logits of the
 * last prompt position. */
int llmc_prefill(llmc_engine* e, int32_t slot, const int32_t* tokens, int32_t n, int32_t* next_token,
                 float* logits_last);

/* Run nsteps greedy decode steps for nslots sequences together.  Each step feeds every
 * slot's pending token, appends it to the cache, and makes the argmax (lowest index on
 * ties) the new pending token.  out_tokens[i*nsteps + s] is the token GENERATED for slot i
 * at step s (the argmax of that step's logits); a provider streams llmc_prefill's
 * *next_token first and then these.  logits (may be NULL) receives [nslots][vocab] of the
 * LAST step.  nslots <= max_slots and nsteps <= 4096 per call; up to 128 sequences share one pass over the
 * weights per step, larger batches run as consecutive groups of 128. */
int llmc_decode(llmc_engine* e, const int32_t* slots, int32_t nslots, int32_t nsteps, int32_t* out_tokens,
                float* logits);

/* Override / read the pending token of a slot (teacher forcing in the parity tests). */
int llmc_seq_set_next_token(llmc_engine* e, int32_t slot, int32_t token);
int llmc_seq_get_next_token(llmc_engine* e, int32_t slot, int32_t* token);

int llmc_engine_stats(llmc_engine* e, llmc_stats* out);
const char* llmc_last_error(llmc_engine* e);
const char* llmc_version(void);

#ifdef __cplusplus
}
#endif
#endif /* LLMC_H */
