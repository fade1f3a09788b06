// decode_wide.cu -- glue kernels of the wide-batch decode step (9..128 sequences per pass over the weights).
//
// With more than 8 rows the matrix products of a decode step are skinny GEMMs: they run on the tcgen05
// kernel of the prefill (gemm_tcgen05.cu: the batch is the M tile, weights streamed once by TMA, activations
// as bf16 hi+lo pairs) and stay HBM-bound.  What differs from the prefill is the bookkeeping around them,
// because the rows of the batch are B different sequences, each at its own position:
//   embed_slots            x[b] = embedding row of next_token[slots[b]]
//   rope_append_batch      RoPE at position seq_len[slot] of each sequence, K/V appended to that sequence's page
//   split_rows             fp32 attention output -> bf16 hi+lo GEMM operand
//                          (both now folded into attn_decode_kernel on the engine's path; kept for the comparator)
//   argmax_finalise        greedy argmax per row (lowest index on ties), next_token / seq_len / output token
// All HBM-bound and negligible next to the weight stream.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include "common.cuh"
#include "kernels.h"

namespace llmc {

namespace {

__global__ void __launch_bounds__(256) embed_slots_kernel(const uint16_t* __restrict__ emb,
                                                          const int* __restrict__ next_token,
                                                          const int* __restrict__ slots, int H, float* __restrict__ x) {
  const int b = blockIdx.x;
  const int tok = next_token[slots[b]];
  const uint16_t* e = emb + (size_t)tok * H;
  float* xr = x + (size_t)b * H;
  for (int k = threadIdx.x * 8; k < H; k += blockDim.x * 8) {
    const uint4 w = *reinterpret_cast<const uint4*>(e + k);
    *reinterpret_cast<float4*>(xr + k) = make_float4(bf16lo(w.x), bf16hi(w.x), bf16lo(w.y), bf16hi(w.y));
    *reinterpret_cast<float4*>(xr + k + 4) = make_float4(bf16lo(w.z), bf16hi(w.z), bf16lo(w.w), bf16hi(w.w));
  }
}

// grid (B, heads); thread i < d/2 rotates the pair (i, i + d/2) of one head (same arithmetic as rope_append_kernel)
__global__ void rope_append_batch_kernel(const float* __restrict__ qkv, const int* __restrict__ slots,
                                         const int* __restrict__ seq_len, const float* __restrict__ rope_cos,
                                         const float* __restrict__ rope_sin, const int* __restrict__ block_table,
                                         int bt_stride, void* __restrict__ kpool, void* __restrict__ vpool, int kv_fp32,
                                         int page_size, int q_heads, int kv_heads, int d, float* __restrict__ q_out) {
  const int b = blockIdx.x, h = blockIdx.y, i = threadIdx.x;
  const int hd2 = d >> 1;
  pdl_launch_dependents();
  pdl_wait();
  if (i >= hd2) return;
  const int slot = slots[b];
  const int pos = seq_len[slot];
  const int rows = (q_heads + 2 * kv_heads) * d;
  const float* src = qkv + (size_t)b * rows + (size_t)h * d;
  float v0 = src[i], v1 = src[i + hd2];
  if (h < q_heads + kv_heads) {
    const float c = rope_cos[(size_t)pos * hd2 + i], s = rope_sin[(size_t)pos * hd2 + i];
    const float o0 = v0 * c - v1 * s, o1 = v1 * c + v0 * s;
    v0 = o0;
    v1 = o1;
  }
  if (h < q_heads) {
    float* qp = q_out + (size_t)b * q_heads * d + (size_t)h * d + i;
    qp[0] = v0;
    qp[hd2] = v1;
  } else {
    const bool isk = h < q_heads + kv_heads;
    const int kvh = isk ? h - q_heads : h - q_heads - kv_heads;
    const int page = block_table[(size_t)slot * bt_stride + pos / page_size];
    const int off = pos % page_size;
    const size_t eo = (((size_t)page * kv_heads + kvh) * page_size + off) * d + i;
    if (kv_fp32) {
      float* dst = reinterpret_cast<float*>(isk ? kpool : vpool) + eo;
      dst[0] = v0;
      dst[hd2] = v1;
    } else {
      uint16_t* dst = reinterpret_cast<uint16_t*>(isk ? kpool : vpool) + eo;
      dst[0] = f32_to_bf16_bits(v0);
      dst[hd2] = f32_to_bf16_bits(v1);
    }
  }
}

__global__ void __launch_bounds__(256) split_rows_kernel(const float* __restrict__ x, size_t n4,
                                                         uint16_t* __restrict__ hi, uint16_t* __restrict__ lo) {
  pdl_launch_dependents();
  pdl_wait();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(x + 4 * i);
    const float y[4] = {v.x, v.y, v.z, v.w};
    uint16_t h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = f32_to_bf16_bits(y[j]);
      l[j] = f32_to_bf16_bits(y[j] - __uint_as_float((uint32_t)h[j] << 16));
    }
    *reinterpret_cast<uint2*>(hi + 4 * i) = make_uint2((uint32_t)h[0] | ((uint32_t)h[1] << 16), (uint32_t)h[2] | ((uint32_t)h[3] << 16));
    if (lo)
      *reinterpret_cast<uint2*>(lo + 4 * i) = make_uint2((uint32_t)l[0] | ((uint32_t)l[1] << 16), (uint32_t)l[2] | ((uint32_t)l[3] << 16));
  }
}

// Greedy argmax in two stages.  Stage 1: grid (chunks, B), each block scans one chunk of a row and leaves a
// (value, index) candidate.  Stage 2: one warp per row reduces the candidates (lowest index on ties), writes
// next_token / seq_len / the output token; the last row to finish (ticket) advances the step counter.
constexpr int AM_CHUNKS = 32;

__global__ void __launch_bounds__(256) argmax_partial_kernel(const float* __restrict__ logits, int V,
                                                             float* __restrict__ part_val, int* __restrict__ part_idx) {
  __shared__ float s_val[8];
  __shared__ int s_idx[8];
  const int chunk = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int per = (V + AM_CHUNKS - 1) / AM_CHUNKS;
  const int lo = chunk * per, hi = min(V, lo + per);
  const float* row = logits + (size_t)b * V;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = lo + tid; i < hi; i += 256) {   // ascending i per thread: strict > keeps the lowest index
    const float v = row[i];
    if (v > best) { best = v; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) { s_val[warp] = best; s_idx[warp] = bi; }
  __syncthreads();
  if (tid == 0) {
#pragma unroll
    for (int w = 1; w < 8; ++w)
      if (s_val[w] > best || (s_val[w] == best && s_idx[w] < bi)) { best = s_val[w]; bi = s_idx[w]; }
    part_val[b * AM_CHUNKS + chunk] = best;
    part_idx[b * AM_CHUNKS + chunk] = bi;
  }
}

__global__ void __launch_bounds__(32) argmax_finalise_kernel(const float* __restrict__ part_val,
                                                             const int* __restrict__ part_idx,
                                                             const int* __restrict__ slots, int* __restrict__ next_token,
                                                             int* __restrict__ seq_len, int* __restrict__ out_tokens,
                                                             int out_stride, int* __restrict__ step_counter,
                                                             unsigned* __restrict__ ticket) {
  const int b = blockIdx.x, lane = threadIdx.x;
  float best = part_val[b * AM_CHUNKS + lane];
  int bi = part_idx[b * AM_CHUNKS + lane];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) {
    const int slot = slots[b];
    const int step = *step_counter;
    next_token[slot] = bi;
    seq_len[slot] += 1;
    out_tokens[(size_t)b * out_stride + step] = bi;
    __threadfence();
    if (atomicAdd(ticket, 1u) == gridDim.x - 1) {   // every row has read the counter
      *step_counter = step + 1;
      *ticket = 0u;
    }
  }
}

}  // namespace

cudaError_t launch_embed_slots(const uint16_t* emb, const int* next_token, const int* slots, int B, int H, float* x,
                               cudaStream_t s) {
  embed_slots_kernel<<<B, 256, 0, s>>>(emb, next_token, slots, H, x);
  return cudaGetLastError();
}

cudaError_t launch_rope_append_batch(const float* qkv, int B, const int* slots, const int* seq_len,
                                     const float* rope_cos, const float* rope_sin, const int* block_table,
                                     int bt_stride, void* kpool, void* vpool, int kv_fp32, int page_size, int q_heads,
                                     int kv_heads, int head_dim, float* q_out, cudaStream_t s, bool pdl) {
  dim3 grid(B, q_heads + 2 * kv_heads);
  return launch_maybe_pdl(rope_append_batch_kernel, grid, dim3(head_dim / 2 < 32 ? 32 : head_dim / 2), 0, s, pdl, qkv,
                          slots, seq_len, rope_cos, rope_sin, block_table, bt_stride, kpool, vpool, kv_fp32, page_size,
                          q_heads, kv_heads, head_dim, q_out);
}

cudaError_t launch_split_rows(const float* x, size_t n, uint16_t* hi, uint16_t* lo, cudaStream_t s, bool pdl) {
  if (n % 4) return cudaErrorInvalidValue;
  const size_t n4 = n / 4;
  const int blocks = (int)std::min<size_t>((n4 + 255) / 256, 1184);
  return launch_maybe_pdl(split_rows_kernel, dim3(blocks), dim3(256), 0, s, pdl, x, n4, hi, lo);
}

size_t argmax_scratch_elems(int B) { return (size_t)B * AM_CHUNKS; }

cudaError_t launch_argmax_finalise(const float* logits, int B, int V, const int* slots, int* next_token, int* seq_len,
                                   int* out_tokens, int out_stride, int* step_counter, unsigned* ticket,
                                   float* part_val, int* part_idx, cudaStream_t s) {
  static_assert(AM_CHUNKS == 32, "stage 2 reduces one candidate per lane");
  argmax_partial_kernel<<<dim3(AM_CHUNKS, B), 256, 0, s>>>(logits, V, part_val, part_idx);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  argmax_finalise_kernel<<<B, 32, 0, s>>>(part_val, part_idx, slots, next_token, seq_len, out_tokens, out_stride,
                                           step_counter, ticket);
  return cudaGetLastError();
}

}  // namespace llmc
