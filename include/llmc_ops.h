/* llmc_ops.h -- kernel-level test and bench hooks exported by libllmc.so.
 *
 * These are NOT part of the provider boundary (that is include/llmc.h).  They let the parity
 * tests drive single kernels of the hot path with host buffers and compare them with the CPU
 * oracle; each call allocates its device buffers, runs, copies back and frees.  They have no
 * counterpart in the reference, whose providers hold no arithmetic
 * (/root/reference/internal/provider/provider.go:13-21).
 *
 * All return 0 or a negative LLMC_ERR_* code; llmc_ops_last_error() has the message.
 * bf16 data crosses the boundary as uint16_t bit patterns.
 */
#ifndef LLMC_OPS_H
#define LLMC_OPS_H

#include <stdint.h>
#include "llmc.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* llmc_ops_last_error(void);

/* Decode GEMV: y[b][n] (+)= sum_k W[n][k] * xhat[b][k], B <= 8, K % 256 == 0, N % 2 == 0.
 * norm != 0: xhat = RMSNorm(x; gain, eps) fused in the prologue; resid != 0: y += (y is in/out). */
int llmc_op_gemv(int device, int N, int K, int B, const uint16_t* W, const float* x, int norm,
                 const uint16_t* gain, float eps, int resid, float* y);

/* Fused gate/up GEMV + SwiGLU on logical matrices Wg, Wu [I][K] (packed/interleaved inside):
 * act[b][j] = silu(Wg xhat)_j * (Wu xhat)_j with xhat = RMSNorm(x; gain, eps). */
int llmc_op_gemv_swiglu(int device, int I, int K, int B, const uint16_t* Wg, const uint16_t* Wu, const float* x,
                        const uint16_t* gain, float eps, float* act);

/* LM head + greedy argmax: logits[b][n] (may be NULL) and tok[b] = lowest-index argmax. */
int llmc_op_lmhead(int device, int V, int K, int B, const uint16_t* W, const float* x, const uint16_t* gain,
                   float eps, float* logits, int32_t* tok);

/* Single-query attention over a paged cache built from dense K/V:
 *   q [B][q_heads*d] fp32; k, v [B][max_ctx][kv_heads*d] bf16 bits (kv_fp32 = 0) or fp32 (1);
 *   ctx[b] tokens valid; out like q.  nsplit = 0 picks the engine's default.  Pages are assigned
 *   in a scrambled order and unwritten slots hold NaN patterns. */
int llmc_op_attn_decode(int device, int B, int q_heads, int kv_heads, int head_dim, int page_size, int nsplit,
                        int max_ctx, int kv_fp32, const int32_t* ctx, const float* q, const void* k, const void* v,
                        float* out);

/* Prefill GEMM C = (A + A_lo) W^T.  A_lo may be NULL (single bf16 operand).  impl 0: tcgen05;
 * 1: plain CUDA-core comparator; 2: tcgen05 with split-K allowed;
 * 3: the decode kernel (operands swapped: weights as the M tile, T <= 128; epi 1, 2, 3 only).  epi: 0 bf16 store, 1 f32 store, 2 f32 residual add (out is
 * in/out), 3 SwiGLU on logical Wg|Wu given as W = [Wg ; Wu] ([2I][K], packed inside), out bf16
 * [T][I] with the optional low half in out_lo (NULL otherwise). */
int llmc_op_gemm(int device, int impl, int epi, int T, int N, int K, const uint16_t* A, const uint16_t* A_lo,
                 const uint16_t* W, void* out, void* out_lo);

/* Times the prefill GEMM on device-resident operands (CUDA events around `iters` launches, 3 warm-up
 * launches first); split != 0 uses the hi+lo A operand.  For the tensor-pipe utilisation figures. */
int llmc_op_gemm_bench(int device, int impl, int epi, int split, int T, int N, int K, int iters, float* ms_per_call);

/* Row-wise RMSNorm fp32 -> bf16 (hi, and lo if out_lo != NULL: x = hi + lo). */
int llmc_op_rmsnorm_rows(int device, int T, int H, const float* x, const uint16_t* gain, float eps, uint16_t* out,
                         uint16_t* out_lo);

/* RoPE + KV append + causal attention of T new rows at positions p0..p0+T-1:
 *   qkv [T][(nq+2nkv)*d] fp32 (pre-RoPE) for the new rows; k_past, v_past [p0][nkv*d] (already
 *   rotated) for the cached prefix, bf16 bits or fp32 per kv_fp32; out (+ out_lo, may be NULL)
 *   bf16 [T][nq*d]; q_out fp32 [T][nq*d]; k_new, v_new [T][nkv*d] read back from the pages.
 *   impl 0: the engine's choice (tensor-core kernel for head_dim 64/128, GQA group 1/2/4, page_size % 64 == 0,
 *   else the CUDA-core kernel); impl 1: always the CUDA-core kernel. */
int llmc_op_prefill_attention(int device, int T, int p0, int q_heads, int kv_heads, int head_dim, int page_size,
                              float rope_theta, int kv_fp32, const float* qkv, const void* k_past, const void* v_past,
                              uint16_t* out, uint16_t* out_lo, float* q_out, void* k_new, void* v_new, int impl);

/* Read a LOGICAL weight tensor of an engine back (un-packing the HBM layout); kinds as in the
 * oracle: layer -1: 0 embedding, 1 lm head, 2 final gain; layer >= 0: 0 g1, 1 wq, 2 wk, 3 wv,
 * 4 wo, 5 g2, 6 wg, 7 wu, 8 wd.  Returns the element count (or < 0); copies if cap suffices. */
int64_t llmc_debug_tensor(llmc_engine* e, int layer, int kind, uint16_t* out, int64_t cap);

/* Read the cached K and V rows ([kv_heads*d] each, in the engine's KV element type) of one
 * position of a slot. */
int llmc_debug_kv(llmc_engine* e, int slot, int layer, int pos, void* k, void* v);

/* LM head + greedy argmax WITHOUT any device allocation.  No counterpart in the reference (it has no
 * device code at all); it exists for callers that must not allocate: processes whose GPU memory is
 * handed out up front by their own arena, plugins inside a framework that owns all allocations, and
 * latency-critical loops where cudaMalloc/cudaFree (device-wide synchronisation, driver locks) are
 * banned after start-up.
 *
 *   logits[b][n] = sum_k W[n][k] * RMSNorm(x[b]; gain, eps)[k]     (d_logits may be NULL)
 *   d_tokens[b]  = lowest-index argmax of row b
 *
 * Every pointer is a DEVICE pointer the caller already owns (W bf16 bits [V][K], x fp32 [B][K], gain
 * bf16 bits [K], logits fp32 [B][V], tokens int32 [B]); `stream` is the caller's cudaStream_t (NULL
 * = default stream).  The call only enqueues work: results are valid once the stream has been
 * synchronised.  The kernel's scratch (per-block argmax candidates, a block counter) sits in a
 * 2 MiB + 512 B array inside the library image, which every GPU that loads the library keeps until
 * exit.  Limits: 1 <= B <= LLMC_NOALLOC_MAX_BATCH, V a multiple of 4 and <= LLMC_NOALLOC_MAX_VOCAB,
 * K a multiple of 256; anything larger is refused with LLMC_ERR_LIMIT.  The first call on a device
 * looks the array's device address up and fills a small table in it; make that first call before any
 * stream capture begins.  All calls share the one array: it is cleared where the kernel needs zeros,
 * but two calls whose kernels can overlap on one GPU (different streams, different threads) race.
 * Neither cudaMalloc/cudaFree nor any allocating library routine is called.  The caller's current
 * device is left as it was (the call selects `device` internally and puts the previous one back). */
#define LLMC_NOALLOC_MAX_BATCH 8
#define LLMC_NOALLOC_MAX_VOCAB 262144
int llmc_lmhead_argmax_noalloc(int device, int V, int K, int B, const uint16_t* d_W, const float* d_x,
                               const uint16_t* d_gain, float eps, float* d_logits, int32_t* d_tokens,
                               void* stream);
const char* llmc_noalloc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LLMC_OPS_H */
