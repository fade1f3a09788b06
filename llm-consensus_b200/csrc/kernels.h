// kernels.h -- launch interfaces of the sm_100a kernels (internal to libllmc.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace llmc {

// ------------------------------------------------------------------ init.cu
cudaError_t launch_init_matrix(uint16_t* dst, int64_t rows, int cols, int64_t ld, uint64_t seed, uint32_t tid,
                               float amp, int mode, int64_t row_off, cudaStream_t s);
cudaError_t launch_init_gain(uint16_t* dst, int n, uint64_t seed, uint32_t tid, cudaStream_t s);

// ------------------------------------------------------------------ gemv.cu (decode)
enum GemvPro { PRO_PLAIN = 0, PRO_NORM = 1, PRO_EMBED_NORM = 2 };
enum GemvEpi { EPI_STORE = 0, EPI_RESID = 1, EPI_QKV_ROPE = 2, EPI_SWIGLU = 3, EPI_LMHEAD = 4 };

constexpr int MAX_GEMV_BATCH = 8;
constexpr int MAX_WIDE_BATCH = 128;  // sequences per pass over the weights on the wide (GEMM) decode path

struct GemvArgs {
  // weights, packed [N][K] bf16, row-major, K % 256 == 0
  const uint16_t* W;
  int N, K;
  int B;  // live batch rows
  // ---- prologue
  const float* x;         // [B][x_stride] fp32 activations (PRO_PLAIN, PRO_NORM)
  int64_t x_stride;
  const uint16_t* gain;   // [K] bf16 RMSNorm gain (PRO_NORM, PRO_EMBED_NORM)
  float eps;
  const uint16_t* emb;    // [V][K] (PRO_EMBED_NORM): x[b] = emb[next_token[slots[b]]]
  const int* slots;       // [B] sequence slot of batch row b
  const int* next_token;  // [max_slots]
  float* x_out;           // [B][K] residual stream written by block 0 (PRO_EMBED_NORM)
  // ---- epilogue
  float* y;               // STORE: y[b][n]; RESID: y[b][n] += ; SWIGLU: act[b][j]; LMHEAD: logits (nullable)
  int64_t y_stride;
  // QKV_ROPE
  const float* rope_cos;  // [max_ctx][head_dim/2]
  const float* rope_sin;
  const int* seq_len;     // [max_slots] tokens already in the cache = position of the new token
  const int* block_table; // [max_slots][bt_stride] page ids
  int bt_stride;
  void* kpool;            // this layer's K pages [page][kv_head][page_size][head_dim], bf16 or fp32
  void* vpool;
  int kv_fp32;
  int page_size, q_heads, kv_heads, head_dim;
  // LMHEAD (greedy argmax, lowest index wins ties; last block finalises the step)
  float* part_val;        // [B][gridDim.x]
  int* part_idx;
  unsigned* counter;      // zero on entry, left zero on exit
  int* next_token_out;    // [max_slots]
  int* seq_len_out;       // [max_slots], += seq_inc
  int seq_inc;
  int* out_tokens;        // [B][out_stride], written at column *step_counter
  int out_stride;
  int* step_counter;      // incremented once per launch
};

// Launch one fused GEMV.  `bt` is the compiled batch tile (1, 2, 4 or 8 >= args.B).
cudaError_t launch_gemv(int epi, int pro, int bt, const GemvArgs& a, bool pdl, cudaStream_t s);
// upper bound of blocks launch_gemv uses for an LM head of N rows (sizes part_val/part_idx)
int gemv_lmhead_blocks(int N);

// gemv_mma.cu: the same fused operation for 2..8 batch rows on the tensor cores (mma.sync tf32, batch = N)
cudaError_t launch_gemv_mma(int epi, int pro, const GemvArgs& a, bool pdl, cudaStream_t s);
int gemv_mma_lmhead_blocks(int N);
// dispatcher used by the engine and the kernel hooks: 1 row -> CUDA-core stream, 2..8 rows -> tensor cores
inline cudaError_t launch_gemv_any(int epi, int pro, const GemvArgs& a, bool pdl, cudaStream_t s) {
  return a.B == 1 ? launch_gemv(epi, pro, 1, a, pdl, s) : launch_gemv_mma(epi, pro, a, pdl, s);
}

// ------------------------------------------------------------------ attn_decode.cu
struct AttnDecodeArgs {
  const float* q;          // [B][q_heads*head_dim] fp32 (RoPE applied); unused when qkv_raw != NULL
  const float* qkv_raw;    // optional: raw QKV rows [B][(q_heads+2kv_heads)*head_dim]; the kernel then applies RoPE to q
  const float *rope_cos, *rope_sin;  //   and appends the rotated K row and the V row itself (pools are written)
  float* out;              // [B][q_heads*head_dim] fp32 (used when out_hi == NULL)
  uint16_t* out_hi;        // optional: write the result as bf16 hi (+ lo if out_lo != NULL) instead of fp32
  uint16_t* out_lo;
  const void* kpool;       // this layer's pages [page][kv_head][page_size][head_dim], bf16 or fp32
  const void* vpool;
  int kv_fp32;
  const int* slots;        // [B]
  const int* seq_len;      // [max_slots]; context length used = seq_len + ctx_extra
  int ctx_extra;           // 1 during decode (the token appended by this step's QKV kernel)
  const int* block_table;
  int bt_stride;
  int page_size, q_heads, kv_heads, head_dim;
  int B, nsplit;
  float* scratch;          // [B][q_heads][nsplit][head_dim + 2] split partials (o, m, l)
  unsigned* counters;      // [B][kv_heads], zero on entry, left zero on exit
};
cudaError_t launch_attn_decode(const AttnDecodeArgs& a, bool pdl, cudaStream_t s);
size_t attn_decode_scratch_floats(int B, int q_heads, int nsplit, int head_dim);
constexpr int ATTN_MAX_SPLIT = 8;
// blocks per (sequence, kv head) for a batch whose longest context is max_ctx_tokens
int attn_decode_nsplit(int max_ctx_tokens);

// ------------------------------------------------------------------ decode_mega.cu (batch-1 decode, one kernel)
struct MegaLayer {
  const uint16_t *wqkv, *wo, *wgu, *wd, *g1, *g2;
  void *kp, *vp;  // this layer's K / V pages
};
struct MegaArgs {
  const MegaLayer* layers;  // device array [L]
  int L, H, QD, KD, I, V, q_heads, kv_heads, head_dim;
  float eps;
  const uint16_t *emb, *lmhead, *gf;
  const float *rope_cos, *rope_sin;
  const int* block_table;
  int slot, bt_stride, page_size, kv_fp32;
  int* seq_len;       // [max_slots]: read at start, written at the end
  int* next_token;    // [max_slots]
  float *x, *q, *attn, *act;   // batch-1 work vectors
  float* logits;      // nullable: logits of the LAST step
  float* attn_scratch;         // [kv_heads][max_tiles][G][head_dim + 2]
  int max_tiles;
  unsigned* attn_counters;     // [kv_heads], zero on entry, left zero
  float* part_val;             // [grid]
  int* part_idx;
  int* out_tokens;             // [nsteps]
  int nsteps;
  unsigned* barrier;           // zero on entry
  unsigned long long* trace;   // nullable: per-barrier (arrive, open) times of block 0, ns
  int attn_tiles_per_block;    // attention: tiles per (kv head, split) block; splits = ceil(tiles / this), <= 8
  int smem_x;                  // filled by the launcher
  int ring_poll_mode;          // decode_ring grid barrier: 1 (default) relaxed spin + one acquire load, 0 acquire-load spin
  // decode_flow.cu only: activations cross SMs as (value, tag) words, no grid barrier
  unsigned long long* tw;      // decode_flow_words() 8-byte words, zero before the first launch
  unsigned tag0;               // first tag of this launch (>= 1); the launch uses nsteps * (5 L + 1) consecutive tags
  int flow_poll_mode;          // 0: one lane per warp watches the first word, then the warp reads; 1: every thread polls
  // decode_ring_batch.cu only: B = 2..4 sequences; x / q / attn / act are [B] rows, logits [B][V], part_val / part_idx
  // [4][grid], attn_counters [4][kv_heads], attn_scratch decode_ring_batch_scratch_floats()
  const int* slots;            // device [B]
  int B;
  int out_stride;              // out_tokens[seq * out_stride + step]
  // attention output and SwiGLU activation of the B sequences as bf16 hi / mid planes ([B][QD], [B][I]): written once by
  // the producing phase, read by every block of the next one without any conversion
  uint16_t *attn_hi, *attn_lo, *act_hi, *act_lo;
  uint16_t *attn_lo2, *act_lo2;   // third plane (hi + mid + lo = 24 mantissa bits): the single-sequence variant  // the residual stream after each projection, already multiplied by the NEXT RMSNorm's gain, as bf16 hi / mid planes
  // [B][H], and its sum of squares per sequence: 64-bit fixed-point accumulators (2^-24), 4 slots x 4 sequences used in
  // turn, zero at launch
  uint16_t *xn_hi, *xn_lo;
  unsigned long long* ss_acc;
};
cudaError_t launch_decode_mega(const MegaArgs& a, int sm_count, cudaStream_t s);
// ------------------------------------------------------------------ decode_ring.cu
// same contract as launch_decode_mega; weights travel through a per-SM shared-memory ring filled by TMA bulk copies
bool decode_ring_supported(int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim);
cudaError_t launch_decode_ring(const MegaArgs& a, int sm_count, cudaStream_t s);
// ------------------------------------------------------------------ decode_ring_mma.cu
// the ring kernel with the slot arithmetic on mma.sync (weights = B operand, activation = hi / mid / lo rows of A)
bool decode_ring_mma_supported(int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim);
cudaError_t launch_decode_ring_mma(const MegaArgs& a, int sm_count, cudaStream_t s);
// ------------------------------------------------------------------ decode_ring_batch.cu
// the mma.sync ring kernel for 2..4 sequences per launch (pairs of sequences as the B columns, hi + mid parts)
bool decode_ring_batch_supported(int B, int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim,
                                 int sm_count);
size_t decode_ring_batch_scratch_floats(int q_heads, int head_dim);
cudaError_t launch_decode_ring_batch(const MegaArgs& a, int sm_count, cudaStream_t s);
// ------------------------------------------------------------------ decode_flow.cu
// the ring kernel without grid barriers: every activation travels as an 8-byte (value, tag) word that its readers
// poll, so a phase hand-over costs one store-to-load trip through L2 instead of fence + arrive + poll + reload
bool decode_flow_supported(int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim, int sm_count);
size_t decode_flow_words(int H, int QD, int KD, int I, int q_heads, int kv_heads, int head_dim, int sm_count);
cudaError_t launch_decode_flow(const MegaArgs& a, int sm_count, cudaStream_t s);

// ------------------------------------------------------------------ prefill.cu
cudaError_t launch_embed_rows(const uint16_t* emb, const int* tokens, int T, int H, float* x, cudaStream_t s);
// RMSNorm of fp32 rows -> bf16 GEMM operand.  out_lo != NULL: split form, x = hi + lo with both bf16
// (the pair carries 16 mantissa bits; the GEMM then accumulates A_hi W + A_lo W).
cudaError_t launch_rmsnorm_rows(const float* x, const uint16_t* gain, float eps, int T, int H, uint16_t* out_hi,
                                uint16_t* out_lo, cudaStream_t s, bool pdl = false, float* zero_rows = nullptr,
                                int zero_cols = 0);
// qkv fp32 [T][qd+2kd] -> q fp32 [T][qd] (RoPE), k/v appended to the pages at positions p0..p0+T-1
cudaError_t launch_rope_append(const float* qkv, int T, int p0, int slot, const float* rope_cos,
                               const float* rope_sin, const int* block_table, int bt_stride, void* kpool,
                               void* vpool, int kv_fp32, int page_size, int q_heads, int kv_heads, int head_dim,
                               float* q_out, cudaStream_t s);
// causal attention of T new rows (positions p0..p0+T-1) over the paged cache; out bf16 [T][qd] (hi, lo).
// impl 0: tensor-core kernel whenever the shape allows it and a workspace is given -- tcgen05 (prefill_attn_tc.cu) for
// head sizes 64 / 128, else mma.sync (prefill_attn_mma.cu) -- else the CUDA-core kernel; impl 1: always the CUDA-core
// kernel (comparator); impl 2: the mma.sync kernel (comparator).
cudaError_t launch_prefill_attention(const float* q, int T, int p0, int slot, const int* block_table,
                                     int bt_stride, const void* kpool, const void* vpool, int kv_fp32, int page_size,
                                     int q_heads, int kv_heads, int head_dim, uint16_t* out_hi, uint16_t* out_lo,
                                     uint16_t* mma_workspace, int impl, cudaStream_t s);
// ------------------------------------------------------------------ prefill_attn_mma.cu
bool prefill_attn_mma_supported(int head_dim, int G);
// split-KV workspace for sequences of up to max_positions cached positions (one layer at a time)
size_t prefill_attn_mma_workspace_bytes(int max_positions, int kv_heads, int head_dim);
cudaError_t launch_prefill_attention_mma(const float* q, int T, int p0, int slot, const int* block_table,
                                         int bt_stride, const void* kpool, const void* vpool, int kv_fp32,
                                         int page_size, int q_heads, int kv_heads, int head_dim, uint16_t* out_hi,
                                         uint16_t* out_lo, uint16_t* workspace, cudaStream_t s);

// ------------------------------------------------------------------ prefill_attn_tc.cu
// the same attention on the 5th-generation tensor cores (tcgen05.mma, S / PV accumulators in TMEM, K and V^T tiles by TMA);
// same arguments and workspace size as the mma.sync version (the workspace layout differs: V planes are transposed)
bool prefill_attn_tc_supported(int head_dim, int G);
cudaError_t launch_prefill_attention_tc(const float* q, int T, int p0, int slot, const int* block_table,
                                        int bt_stride, const void* kpool, const void* vpool, int kv_fp32,
                                        int page_size, int q_heads, int kv_heads, int head_dim, uint16_t* out_hi,
                                        uint16_t* out_lo, uint16_t* workspace, cudaStream_t s);

// ------------------------------------------------------------------ decode_wide.cu (decode, 9..128 sequences)
cudaError_t launch_embed_slots(const uint16_t* emb, const int* next_token, const int* slots, int B, int H, float* x,
                               cudaStream_t s);
cudaError_t launch_rope_append_batch(const float* qkv, int B, const int* slots, const int* seq_len,
                                     const float* rope_cos, const float* rope_sin, const int* block_table,
                                     int bt_stride, void* kpool, void* vpool, int kv_fp32, int page_size, int q_heads,
                                     int kv_heads, int head_dim, float* q_out, cudaStream_t s, bool pdl = false);
cudaError_t launch_split_rows(const float* x, size_t n, uint16_t* hi, uint16_t* lo, cudaStream_t s, bool pdl = false);
// greedy argmax of B logit rows (lowest index on ties); writes next_token / seq_len (+1) / out_tokens[b][*step_counter]
// and advances the step counter once per launch
size_t argmax_scratch_elems(int B);   // part_val / part_idx sizes
cudaError_t launch_argmax_finalise(const float* logits, int B, int V, const int* slots, int* next_token, int* seq_len,
                                   int* out_tokens, int out_stride, int* step_counter, unsigned* ticket,
                                   float* part_val, int* part_idx, cudaStream_t s);

// ------------------------------------------------------------------ gemm_tcgen05.cu / gemm_simple.cu (prefill)
// C[T][N] = (A_hi [+ A_lo])[T][K] (bf16) * W[N][K]^T (bf16), fp32 accumulation.  A_lo may be NULL.
enum GemmEpi {
  GEMM_STORE_BF16 = 0,  // out bf16 [T][N]
  GEMM_STORE_F32 = 1,   // out f32  [T][N]
  GEMM_RESID_F32 = 2,   // out f32  [T][N] += acc
  GEMM_SWIGLU_BF16 = 3  // W rows gate/up interleaved in blocks of 64; out bf16 [T][N/2] = silu(g)*u (hi [, lo])
};
// allow_ksplit: decode-shaped calls (T <= 128, fp32 output) may split K over several CTAs per output tile, which add
// their partial sums with fp32 reductions (summation order then varies from run to run in the last bits)
cudaError_t launch_gemm_tcgen05(const uint16_t* A_hi, const uint16_t* A_lo, const uint16_t* W, void* out,
                                void* out_lo, int T, int N, int K, int epi, cudaStream_t s, bool allow_ksplit = false,
                                bool pdl = false);
// ------------------------------------------------------------------ gemm_decode_tc.cu (decode, <= 128 sequences)
// Same product with the operands swapped on the tensor cores (weights = M tile, batch = N): out[B][N] fp32 store /
// residual add / SwiGLU (bf16 hi [, lo]); split-K where a matrix has few row tiles.
bool gemm_decode_tc_supported(int B, int N, int K, int epi);
cudaError_t launch_gemm_decode_tc(const uint16_t* A_hi, const uint16_t* A_lo, const uint16_t* W, void* out,
                                  void* out_lo, int B, int N, int K, int epi, cudaStream_t s, bool pdl = false);
cudaError_t launch_gemm_simple(const uint16_t* A_hi, const uint16_t* A_lo, const uint16_t* W, void* out,
                               void* out_lo, int T, int N, int K, int epi, cudaStream_t s);

}  // namespace llmc
