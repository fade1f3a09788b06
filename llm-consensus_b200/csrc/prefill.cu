// prefill.cu -- the row-wise kernels around the prefill GEMMs: embedding gather, RMSNorm
// (fp32 residual stream -> bf16 GEMM operand, optionally as a hi+lo bf16 pair), RoPE + paged-KV
// append (bf16 or fp32 cache), causal attention.
//
// All are HBM/L2-bound elementwise or reduction kernels except the attention, which is the
// first (CUDA-core, shared-memory tiled) version: one warp per query row, the q heads of a
// GQA group served from one staged K/V tile.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the
// boundary; the providers behind it are remote HTTP calls).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"
#include "attn_common.cuh"

namespace llmc {

namespace {

__global__ void __launch_bounds__(256) embed_rows_kernel(const uint16_t* __restrict__ emb, const int* __restrict__ tokens,
                                                         int T, int H, float* __restrict__ x) {
  const int t = blockIdx.x;
  const int tok = tokens[t];
  const uint16_t* e = emb + (size_t)tok * H;
  float* xr = x + (size_t)t * H;
  for (int k = threadIdx.x * 8; k < H; k += blockDim.x * 8) {
    uint4 w = *reinterpret_cast<const uint4*>(e + k);
    *reinterpret_cast<float4*>(xr + k) = make_float4(bf16lo(w.x), bf16hi(w.x), bf16lo(w.y), bf16hi(w.y));
    *reinterpret_cast<float4*>(xr + k + 4) = make_float4(bf16lo(w.z), bf16hi(w.z), bf16lo(w.w), bf16hi(w.w));
  }
}

// v = hi + lo with both bf16: the pair keeps 16 mantissa bits of the fp32 value
__device__ __forceinline__ void split_bf16(float v, uint16_t& hi, uint16_t& lo) {
  hi = f32_to_bf16_bits(v);
  lo = f32_to_bf16_bits(v - __uint_as_float((uint32_t)hi << 16));
}

// one block (256 threads) per row; two passes over a row that stays in L1/L2

// zero_row != NULL: the block also clears row t (zero_cols floats) of another buffer -- the wide decode step uses
// it to prepare the accumulation target of the following split-K GEMM without a separate memset node
__global__ void __launch_bounds__(256) rmsnorm_rows_kernel(const float* __restrict__ x, const uint16_t* __restrict__ gain,
                                                           float eps, int T, int H, uint16_t* __restrict__ out,
                                                           uint16_t* __restrict__ out_lo, float* __restrict__ zero_row,
                                                           int zero_cols) {
  __shared__ float red[8];
  const int t = blockIdx.x, tid = threadIdx.x;
  const float* xr = x + (size_t)t * H;
  pdl_launch_dependents();
  pdl_wait();
  if (zero_row)
    for (int k = tid * 4; k < zero_cols; k += 256 * 4)
      *reinterpret_cast<float4*>(zero_row + (size_t)t * zero_cols + k) = make_float4(0.f, 0.f, 0.f, 0.f);
  float ss = 0.f;
  for (int k = tid * 4; k < H; k += 256 * 4) {
    float4 v = *reinterpret_cast<const float4*>(xr + k);
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = warp_sum(ss);
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float r = rsqrtf(tot / (float)H + eps);
  for (int k = tid * 4; k < H; k += 256 * 4) {
    float4 v = *reinterpret_cast<const float4*>(xr + k);
    uint2 g = *reinterpret_cast<const uint2*>(gain + k);
    const float y[4] = {v.x * r * bf16lo(g.x), v.y * r * bf16hi(g.x), v.z * r * bf16lo(g.y), v.w * r * bf16hi(g.y)};
    uint16_t hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) split_bf16(y[i], hi[i], lo[i]);
    *reinterpret_cast<uint2*>(out + (size_t)t * H + k) =
        make_uint2((uint32_t)hi[0] | ((uint32_t)hi[1] << 16), (uint32_t)hi[2] | ((uint32_t)hi[3] << 16));
    if (out_lo)
      *reinterpret_cast<uint2*>(out_lo + (size_t)t * H + k) =
          make_uint2((uint32_t)lo[0] | ((uint32_t)lo[1] << 16), (uint32_t)lo[2] | ((uint32_t)lo[3] << 16));
  }
}

// grid (T, heads); thread i < d/2 rotates the pair (i, i + d/2) of one head
__global__ void rope_append_kernel(const float* __restrict__ qkv, int T, int p0, int slot,
                                   const float* __restrict__ rope_cos, const float* __restrict__ rope_sin,
                                   const int* __restrict__ block_table, int bt_stride, void* __restrict__ kpool,
                                   void* __restrict__ vpool, int kv_fp32, int page_size, int q_heads, int kv_heads,
                                   int d, float* __restrict__ q_out) {
  const int t = blockIdx.x, h = blockIdx.y, i = threadIdx.x;
  const int hd2 = d >> 1;
  if (i >= hd2) return;
  const int pos = p0 + t;
  const int rows = (q_heads + 2 * kv_heads) * d;
  const float* src = qkv + (size_t)t * rows + (size_t)h * d;
  float v0 = src[i], v1 = src[i + hd2];
  if (h < q_heads + kv_heads) {
    const float c = rope_cos[(size_t)pos * hd2 + i], s = rope_sin[(size_t)pos * hd2 + i];
    const float o0 = v0 * c - v1 * s, o1 = v1 * c + v0 * s;
    v0 = o0;
    v1 = o1;
  }
  if (h < q_heads) {
    float* qp = q_out + (size_t)t * q_heads * d + (size_t)h * d + i;
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

constexpr int PROWS = 8;  // query rows (= warps) per block

template <int D, int G, typename KVT>
__global__ void __launch_bounds__(PROWS * 32) prefill_attention_kernel(
    const float* __restrict__ q, int T, int p0, int slot, const int* __restrict__ block_table, int bt_stride,
    const KVT* __restrict__ kpool, const KVT* __restrict__ vpool, int page_size, int q_heads, int kv_heads,
    uint16_t* __restrict__ out, uint16_t* __restrict__ out_lo) {
  constexpr int KS = TileStride<KVT, D>::value;
  constexpr int EPV = 16 / (int)sizeof(KVT);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  KVT* ks = reinterpret_cast<KVT*>(smem_raw);
  KVT* vs = ks + ATT_TP * KS;
  float* qs = reinterpret_cast<float*>(vs + ATT_TP * KS);  // [PROWS][G][D]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kvh = blockIdx.y;
  const int t = blockIdx.x * PROWS + warp;
  const bool live = t < T;
  const int QD = q_heads * D;
  const float scale = rsqrtf((float)D);

  float* myq = qs + (size_t)warp * G * D;
  if (live) {
    const float* qr = q + (size_t)t * QD + (size_t)kvh * G * D;
    for (int i = lane; i < G * D; i += 32) myq[i] = qr[i];
  }
  float m[G], l[G], o[G][D / 32];
#pragma unroll
  for (int j = 0; j < G; ++j) {
    m[j] = -INFINITY;
    l[j] = 0.f;
#pragma unroll
    for (int i = 0; i < D / 32; ++i) o[j][i] = 0.f;
  }
  const int tmax = min(T - 1, blockIdx.x * PROWS + PROWS - 1);
  const int ntiles = (p0 + tmax + 1 + ATT_TP - 1) / ATT_TP;
  const int myctx = live ? p0 + t + 1 : 0;

  for (int tile = 0; tile < ntiles; ++tile) {
    const int pbase = tile * ATT_TP;
    const int tile_valid = min(ATT_TP, p0 + tmax + 1 - pbase);
    const int page = block_table[(size_t)slot * bt_stride + pbase / page_size];
    const size_t base = (((size_t)page * kv_heads + kvh) * page_size + (pbase % page_size)) * D;
    __syncthreads();
    for (int c = tid; c < ATT_TP * (D / EPV); c += PROWS * 32) {
      const int r = c / (D / EPV), cc = c % (D / EPV);
      uint4 kk = make_uint4(0, 0, 0, 0), vv = kk;
      if (r < tile_valid) {
        kk = *reinterpret_cast<const uint4*>(kpool + base + (size_t)r * D + cc * EPV);
        vv = *reinterpret_cast<const uint4*>(vpool + base + (size_t)r * D + cc * EPV);
      }
      *reinterpret_cast<uint4*>(ks + r * KS + cc * EPV) = kk;
      *reinterpret_cast<uint4*>(vs + r * KS + cc * EPV) = vv;
    }
    __syncthreads();
    const int valid = min(ATT_TP, myctx - pbase);
    if (valid > 0) {
#pragma unroll
      for (int j = 0; j < G; ++j) attend_tile<D, KVT>(myq + j * D, ks, vs, valid, scale, m[j], l[j], o[j]);
    }
  }
  if (live) {
#pragma unroll
    for (int j = 0; j < G; ++j) {
      const float inv = 1.0f / l[j];
      const size_t off = (size_t)t * QD + (size_t)(kvh * G + j) * D + lane * (D / 32);
#pragma unroll
      for (int i = 0; i < D / 32; ++i) {
        uint16_t hi, lo;
        split_bf16(o[j][i] * inv, hi, lo);
        out[off + i] = hi;
        if (out_lo) out_lo[off + i] = lo;
      }
    }
  }
}

template <int D, int G, typename KVT>
cudaError_t launch_pa(const float* q, int T, int p0, int slot, const int* bt, int bt_stride, const void* kpool,
                      const void* vpool, int page_size, int q_heads, int kv_heads, uint16_t* out, uint16_t* out_lo,
                      cudaStream_t s) {
  const size_t smem = (size_t)2 * ATT_TP * TileStride<KVT, D>::value * sizeof(KVT) + (size_t)PROWS * G * D * 4;
  auto kern = prefill_attention_kernel<D, G, KVT>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  dim3 grid((T + PROWS - 1) / PROWS, kv_heads);
  kern<<<grid, PROWS * 32, smem, s>>>(q, T, p0, slot, bt, bt_stride, reinterpret_cast<const KVT*>(kpool),
                                      reinterpret_cast<const KVT*>(vpool), page_size, q_heads, kv_heads, out, out_lo);
  return cudaGetLastError();
}

template <int D, typename KVT>
cudaError_t launch_pa_g(int G, const float* q, int T, int p0, int slot, const int* bt, int bt_stride,
                        const void* kpool, const void* vpool, int page_size, int q_heads, int kv_heads,
                        uint16_t* out, uint16_t* out_lo, cudaStream_t s) {
  switch (G) {
    case 1: return launch_pa<D, 1, KVT>(q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
    case 2: return launch_pa<D, 2, KVT>(q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
    case 4: return launch_pa<D, 4, KVT>(q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
    case 8: return launch_pa<D, 8, KVT>(q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
  }
  return cudaErrorInvalidValue;
}

template <typename KVT>
cudaError_t launch_pa_d(int D, int G, const float* q, int T, int p0, int slot, const int* bt, int bt_stride,
                        const void* kpool, const void* vpool, int page_size, int q_heads, int kv_heads,
                        uint16_t* out, uint16_t* out_lo, cudaStream_t s) {
  switch (D) {
    case 32: return launch_pa_g<32, KVT>(G, q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
    case 64: return launch_pa_g<64, KVT>(G, q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
    case 128: return launch_pa_g<128, KVT>(G, q, T, p0, slot, bt, bt_stride, kpool, vpool, page_size, q_heads, kv_heads, out, out_lo, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

cudaError_t launch_embed_rows(const uint16_t* emb, const int* tokens, int T, int H, float* x, cudaStream_t s) {
  embed_rows_kernel<<<T, 256, 0, s>>>(emb, tokens, T, H, x);
  return cudaGetLastError();
}

cudaError_t launch_rmsnorm_rows(const float* x, const uint16_t* gain, float eps, int T, int H, uint16_t* out_hi,
                                uint16_t* out_lo, cudaStream_t s, bool pdl, float* zero_rows, int zero_cols) {
  if (zero_rows && zero_cols % 4) return cudaErrorInvalidValue;
  return launch_maybe_pdl(rmsnorm_rows_kernel, dim3(T), dim3(256), 0, s, pdl, x, gain, eps, T, H, out_hi, out_lo,
                          zero_rows, zero_cols);
}

cudaError_t launch_rope_append(const float* qkv, int T, int p0, int slot, const float* rope_cos,
                               const float* rope_sin, const int* block_table, int bt_stride, void* kpool,
                               void* vpool, int kv_fp32, int page_size, int q_heads, int kv_heads, int head_dim,
                               float* q_out, cudaStream_t s) {
  dim3 grid(T, q_heads + 2 * kv_heads);
  rope_append_kernel<<<grid, head_dim / 2 < 32 ? 32 : head_dim / 2, 0, s>>>(
      qkv, T, p0, slot, rope_cos, rope_sin, block_table, bt_stride, kpool, vpool, kv_fp32, page_size, q_heads,
      kv_heads, head_dim, q_out);
  return cudaGetLastError();
}

cudaError_t launch_prefill_attention(const float* q, int T, int p0, int slot, const int* block_table,
                                     int bt_stride, const void* kpool, const void* vpool, int kv_fp32, int page_size,
                                     int q_heads, int kv_heads, int head_dim, uint16_t* out_hi, uint16_t* out_lo,
                                     uint16_t* mma_workspace, int impl, cudaStream_t s) {
  if (q_heads % kv_heads != 0 || page_size % ATT_TP != 0) return cudaErrorInvalidValue;
  const int G = q_heads / kv_heads;
  // tensor-core paths: tcgen05 (prefill_attn_tc.cu) where the shape allows it, else mma.sync (prefill_attn_mma.cu);
  // LLMC_PREFILL_ATTN=mma / tc forces one of them (comparison runs)
  static const char* const pick = getenv("LLMC_PREFILL_ATTN");
  const bool want_tc = pick ? strcmp(pick, "mma") != 0 : true;
  if (impl == 2) impl = 0;   // 2 = the mma.sync kernel (comparator of the tests), otherwise handled like 0
  else if (impl == 0 && mma_workspace && want_tc && prefill_attn_tc_supported(head_dim, G))
    return launch_prefill_attention_tc(q, T, p0, slot, block_table, bt_stride, kpool, vpool, kv_fp32, page_size,
                                       q_heads, kv_heads, head_dim, out_hi, out_lo, mma_workspace, s);
  if (impl == 0 && mma_workspace && prefill_attn_mma_supported(head_dim, G))
    return launch_prefill_attention_mma(q, T, p0, slot, block_table, bt_stride, kpool, vpool, kv_fp32, page_size,
                                        q_heads, kv_heads, head_dim, out_hi, out_lo, mma_workspace, s);
  if (kv_fp32)
    return launch_pa_d<float>(head_dim, G, q, T, p0, slot, block_table, bt_stride, kpool, vpool, page_size, q_heads,
                              kv_heads, out_hi, out_lo, s);
  return launch_pa_d<uint16_t>(head_dim, G, q, T, p0, slot, block_table, bt_stride, kpool, vpool, page_size, q_heads,
                               kv_heads, out_hi, out_lo, s);
}

}  // namespace llmc
