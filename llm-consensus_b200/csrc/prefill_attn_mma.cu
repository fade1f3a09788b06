// prefill_attn_mma.cu -- causal GQA attention for the prefill on the tensor cores (mma.sync bf16, split operands).
//
// Tensor-pipe bound (4 * n_h * d * T^2 / 2 flops per layer, causal).  Flash-attention
// structure: a block of 8 warps owns one kv head and 128 / G query rows -- warp w works on q head (w % G) of the
// GQA group and row tile (w / G) -- so a staged K/V tile serves the whole group; it sweeps the cached positions
// in 64-wide tiles with an online softmax.
//
// Precision: bf16 attention would by itself eat most of the 2e-2 logit budget on this model (DESIGN.md
// section 5), so every operand is carried as a bf16 pair x = hi + mid (16 mantissa bits, the same class as the
// split A operand of the prefill GEMMs) and each product runs as 3 MMAs: hi*hi + mid*hi + hi*mid.
//   * K and V are split once per layer by kv_split_kernel into a dense workspace [kv_head][position][d]
//     (hi and mid planes, zero-padded to a multiple of 64 positions), so the attention kernel streams them
//     with cp.async into a double-buffered shared-memory ring with no conversion in its loop.
//     A bf16 KV cache (LLMC_FLAG_KV_BF16) has no mid part: 2 MMAs per product.
//   * Q is scaled by 1/sqrt(d) and split when the block stages it; P is split in registers: the accumulator
//     layout of S = Q K^T is exactly the A-fragment layout of m16n8k16, so P never leaves the registers.
//   * fragments come from shared memory with ldmatrix (.trans for V); rows are padded by 16 bytes so that
//     every ldmatrix phase is bank-conflict-free.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include "common.cuh"
#include "kernels.h"

namespace llmc {

namespace {

constexpr int KT = 64;      // positions per K/V tile
constexpr int NWARP = 8;

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {   // lo -> bits 0..15
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// (x0, x1) -> packed hi pair and packed mid pair (mid = bf16(x - hi))
__device__ __forceinline__ void split_pair(float x0, float x1, uint32_t& h, uint32_t& m) {
  h = pack_bf16x2(x0, x1);
  m = pack_bf16x2(x0 - __uint_as_float(h << 16), x1 - __uint_as_float(h & 0xffff0000u));
}
__device__ __forceinline__ void mma16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---- K/V pages of one slot -> dense split workspace.  ws planes: khi, kmid, vhi, vmid, each [kv_heads][npad][D].
template <typename KVT>
__global__ void kv_split_kernel(const KVT* __restrict__ kpool, const KVT* __restrict__ vpool, int slot,
                                const int* __restrict__ block_table, int bt_stride, int page_size, int kv_heads,
                                int D, int npos, int npad, uint16_t* __restrict__ ws) {
  const int pos = blockIdx.x, h = blockIdx.y;
  const size_t plane = (size_t)kv_heads * npad * D;
  const size_t dst = ((size_t)h * npad + pos) * D;
  uint16_t* khi = ws + dst;
  uint16_t* kmid = khi + plane;
  uint16_t* vhi = kmid + plane;
  uint16_t* vmid = vhi + plane;
  const int c = threadIdx.x * 4;   // blockDim = D / 4
  if (pos >= npos) {
    const uint2 z = make_uint2(0u, 0u);
    *reinterpret_cast<uint2*>(khi + c) = z;
    *reinterpret_cast<uint2*>(vhi + c) = z;
    if constexpr (sizeof(KVT) == 4) {
      *reinterpret_cast<uint2*>(kmid + c) = z;
      *reinterpret_cast<uint2*>(vmid + c) = z;
    }
    return;
  }
  const int page = block_table[(size_t)slot * bt_stride + pos / page_size];
  const size_t src = (((size_t)page * kv_heads + h) * page_size + (pos % page_size)) * D + c;
  if constexpr (sizeof(KVT) == 4) {
    const float4 k = *reinterpret_cast<const float4*>(kpool + src);
    const float4 v = *reinterpret_cast<const float4*>(vpool + src);
    uint2 hh, mm;
    split_pair(k.x, k.y, hh.x, mm.x); split_pair(k.z, k.w, hh.y, mm.y);
    *reinterpret_cast<uint2*>(khi + c) = hh;
    *reinterpret_cast<uint2*>(kmid + c) = mm;
    split_pair(v.x, v.y, hh.x, mm.x); split_pair(v.z, v.w, hh.y, mm.y);
    *reinterpret_cast<uint2*>(vhi + c) = hh;
    *reinterpret_cast<uint2*>(vmid + c) = mm;
  } else {
    *reinterpret_cast<uint2*>(khi + c) = *reinterpret_cast<const uint2*>(kpool + src);
    *reinterpret_cast<uint2*>(vhi + c) = *reinterpret_cast<const uint2*>(vpool + src);
  }
}

template <int D, int G, bool MID>
__global__ void __launch_bounds__(NWARP * 32, 1) prefill_attn_mma_kernel(
    const float* __restrict__ q, int T, int p0, int npad, const uint16_t* __restrict__ ws, int q_heads,
    int kv_heads, uint16_t* __restrict__ out, uint16_t* __restrict__ out_lo) {
  constexpr int S = D + 8;                  // padded row stride in bf16 elements (16 bytes)
  constexpr int QROWS = 16 * NWARP;         // staged q rows (all warps): [warp][16]
  constexpr int RT = NWARP / G;             // row tiles per block
  constexpr int NPL = MID ? 4 : 2;          // K/V planes per stage
  constexpr int STAGE = NPL * KT * S;       // bf16 elements per stage
  extern __shared__ __align__(16) uint16_t smem[];
  uint16_t* sq_hi = smem;                   // [QROWS][S]
  uint16_t* sq_mid = sq_hi + QROWS * S;
  uint16_t* skv = sq_mid + QROWS * S;       // [2][NPL][KT][S]: khi, vhi, (kmid, vmid)

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int kvh = blockIdx.y;
  const int rb = gridDim.x - 1 - blockIdx.x;                 // long (late) row blocks are scheduled first
  const int head = kvh * G + (warp % G);
  const int r0 = rb * (16 * RT) + (warp / G) * 16;           // first query row of this warp
  const int QD = q_heads * D;
  const float scale = rsqrtf((float)D);
  const size_t plane = (size_t)kv_heads * npad * D;
  const uint16_t* wsk = ws + (size_t)kvh * npad * D;         // khi of this kv head; + plane: kmid, vhi, vmid

  const int last_row = min(T - 1, rb * (16 * RT) + 16 * RT - 1);
  const int ntiles = (p0 + last_row + 1 + KT - 1) / KT;

  auto issue_tile = [&](int tile, int stage) {
    constexpr int CPR = D / 8;                               // 16-byte chunks per row
    const uint32_t dst_s = (uint32_t)__cvta_generic_to_shared(skv + (size_t)stage * STAGE);
    for (int e = tid; e < NPL * KT * CPR; e += NWARP * 32) {
      const int pl = e / (KT * CPR), rem = e - pl * (KT * CPR);
      const int r = rem / CPR, c = (rem - r * CPR) * 8;
      // smem plane order: khi, vhi, kmid, vmid ; workspace plane order: khi, kmid, vhi, vmid
      const int wpl = (pl == 0) ? 0 : (pl == 1) ? 2 : (pl == 2) ? 1 : 3;
      cp_async16(dst_s + (uint32_t)(((pl * KT + r) * S + c) * 2),
                 wsk + (size_t)wpl * plane + ((size_t)tile * KT + r) * D + c);
    }
    cp_async_commit();
  };

  issue_tile(0, 0);

  // ---- stage this warp's 16 q rows (scaled) as hi / mid
  {
    uint16_t* qh = sq_hi + (size_t)warp * 16 * S;
    uint16_t* qm = sq_mid + (size_t)warp * 16 * S;
    for (int e = lane; e < 16 * (D / 4); e += 32) {
      const int r = e / (D / 4), c4 = (e - r * (D / 4)) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r0 + r < T) v = *reinterpret_cast<const float4*>(q + (size_t)(r0 + r) * QD + (size_t)head * D + c4);
      uint2 hh, mm;
      split_pair(v.x * scale, v.y * scale, hh.x, mm.x);
      split_pair(v.z * scale, v.w * scale, hh.y, mm.y);
      *reinterpret_cast<uint2*>(qh + r * S + c4) = hh;
      *reinterpret_cast<uint2*>(qm + r * S + c4) = mm;
    }
    __syncwarp();
  }

  float o[D / 8][4];
#pragma unroll
  for (int db = 0; db < D / 8; ++db)
#pragma unroll
    for (int i = 0; i < 4; ++i) o[db][i] = 0.f;
  float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};    // rows g and g + 8
  const int lim0 = p0 + r0 + g, lim1 = lim0 + 8;             // last visible position of this lane's two rows

  // ldmatrix lane addresses (bytes, relative to the array base)
  const uint32_t sq_hi_s = (uint32_t)__cvta_generic_to_shared(sq_hi + (size_t)warp * 16 * S);
  const uint32_t sq_mid_s = (uint32_t)__cvta_generic_to_shared(sq_mid + (size_t)warp * 16 * S);
  const uint32_t skv_s = (uint32_t)__cvta_generic_to_shared(skv);
  const uint32_t a_off = (uint32_t)(((lane & 15) * S + (lane >> 4) * 8) * 2);                            // Q: + kk*32
  const uint32_t k_off = (uint32_t)((((lane & 7) + (lane >> 4) * 8) * S + ((lane >> 3) & 1) * 8) * 2);   // K: + nb2*16 rows, + kk*32
  const uint32_t v_off = (uint32_t)((((lane & 7) + ((lane >> 3) & 1) * 8) * S + (lane >> 4) * 8) * 2);   // V: + kb*16 rows, + db2*32

  for (int tile = 0; tile < ntiles; ++tile) {
    cp_async_wait_all();
    __syncthreads();                          // tile landed for everyone; everyone is done with the other stage
    if (tile + 1 < ntiles) issue_tile(tile + 1, (tile + 1) & 1);
    const int pbase = tile * KT;
    if (pbase > p0 + min(T - 1, r0 + 15)) continue;           // nothing visible to this warp's rows (warp-uniform)
    const uint32_t st = skv_s + (uint32_t)((tile & 1) * STAGE * 2);
    const uint32_t skh = st, svh = st + KT * S * 2, skm = st + 2 * KT * S * 2, svm = st + 3 * KT * S * 2;

    // ---- S = Q K^T (16 rows x 64 positions per warp)
    float sc[KT / 8][4];
#pragma unroll
    for (int nb = 0; nb < KT / 8; ++nb)
#pragma unroll
      for (int i = 0; i < 4; ++i) sc[nb][i] = 0.f;
#pragma unroll
    for (int kk = 0; kk < D / 16; ++kk) {
      uint32_t ah[4], am[4];
      ldsm4(ah, sq_hi_s + a_off + kk * 32);
      ldsm4(am, sq_mid_s + a_off + kk * 32);
#pragma unroll
      for (int nb2 = 0; nb2 < KT / 16; ++nb2) {
        uint32_t bh[4];
        ldsm4(bh, skh + k_off + nb2 * (16 * S * 2) + kk * 32);
        mma16(sc[2 * nb2], ah, bh[0], bh[1]);
        mma16(sc[2 * nb2 + 1], ah, bh[2], bh[3]);
        mma16(sc[2 * nb2], am, bh[0], bh[1]);
        mma16(sc[2 * nb2 + 1], am, bh[2], bh[3]);
        if constexpr (MID) {
          uint32_t bm[4];
          ldsm4(bm, skm + k_off + nb2 * (16 * S * 2) + kk * 32);
          mma16(sc[2 * nb2], ah, bm[0], bm[1]);
          mma16(sc[2 * nb2 + 1], ah, bm[2], bm[3]);
        }
      }
    }
    // ---- causal mask + online softmax; accumulator columns of this lane: positions nb*8 + 2t, 2t+1
    if (pbase + KT - 1 > lim0 - g) {          // the tile reaches past the first row of the warp: mask
#pragma unroll
      for (int nb = 0; nb < KT / 8; ++nb) {
        const int pos = pbase + nb * 8 + 2 * t;
        if (pos > lim0) sc[nb][0] = -INFINITY;
        if (pos + 1 > lim0) sc[nb][1] = -INFINITY;
        if (pos > lim1) sc[nb][2] = -INFINITY;
        if (pos + 1 > lim1) sc[nb][3] = -INFINITY;
      }
    }
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nb = 0; nb < KT / 8; ++nb) {
      mx[0] = fmaxf(mx[0], fmaxf(sc[nb][0], sc[nb][1]));
      mx[1] = fmaxf(mx[1], fmaxf(sc[nb][2], sc[nb][3]));
    }
    float corr[2], rsum[2] = {0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 1));
      mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], 2));
      const float mn = fmaxf(m[j], mx[j]);    // position 0 is visible to every row, so mn is finite
      corr[j] = __expf(m[j] - mn);
      m[j] = mn;
    }
#pragma unroll
    for (int nb = 0; nb < KT / 8; ++nb) {
      sc[nb][0] = __expf(sc[nb][0] - m[0]);
      sc[nb][1] = __expf(sc[nb][1] - m[0]);
      sc[nb][2] = __expf(sc[nb][2] - m[1]);
      sc[nb][3] = __expf(sc[nb][3] - m[1]);
      rsum[0] += sc[nb][0] + sc[nb][1];
      rsum[1] += sc[nb][2] + sc[nb][3];
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      rsum[j] += __shfl_xor_sync(0xffffffffu, rsum[j], 1);
      rsum[j] += __shfl_xor_sync(0xffffffffu, rsum[j], 2);
      l[j] = l[j] * corr[j] + rsum[j];
    }
#pragma unroll
    for (int db = 0; db < D / 8; ++db) {
      o[db][0] *= corr[0]; o[db][1] *= corr[0];
      o[db][2] *= corr[1]; o[db][3] *= corr[1];
    }
    // ---- O += P V : the S accumulators of position blocks 2kb, 2kb+1 are the A fragment of k-step kb
#pragma unroll
    for (int kb = 0; kb < KT / 16; ++kb) {
      uint32_t ph[4], pm[4];
      split_pair(sc[2 * kb][0], sc[2 * kb][1], ph[0], pm[0]);
      split_pair(sc[2 * kb][2], sc[2 * kb][3], ph[1], pm[1]);
      split_pair(sc[2 * kb + 1][0], sc[2 * kb + 1][1], ph[2], pm[2]);
      split_pair(sc[2 * kb + 1][2], sc[2 * kb + 1][3], ph[3], pm[3]);
#pragma unroll
      for (int db2 = 0; db2 < D / 16; ++db2) {
        uint32_t bh[4];
        ldsm4t(bh, svh + v_off + kb * (16 * S * 2) + db2 * 32);
        mma16(o[2 * db2], ph, bh[0], bh[1]);
        mma16(o[2 * db2 + 1], ph, bh[2], bh[3]);
        mma16(o[2 * db2], pm, bh[0], bh[1]);
        mma16(o[2 * db2 + 1], pm, bh[2], bh[3]);
        if constexpr (MID) {
          uint32_t bm[4];
          ldsm4t(bm, svm + v_off + kb * (16 * S * 2) + db2 * 32);
          mma16(o[2 * db2], ph, bm[0], bm[1]);
          mma16(o[2 * db2 + 1], ph, bm[2], bm[3]);
        }
      }
    }
  }
  cp_async_wait_all();

  // ---- normalise and store as bf16 hi (+ lo): lane holds dims db*8 + 2t, 2t+1 of rows g and g+8
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int row = r0 + g + 8 * j;
    if (row >= T) continue;
    const float inv = 1.0f / l[j];
    const size_t off = (size_t)row * QD + (size_t)head * D + 2 * t;
#pragma unroll
    for (int db = 0; db < D / 8; ++db) {
      const float v0 = o[db][2 * j] * inv, v1 = o[db][2 * j + 1] * inv;
      const uint16_t h0 = f32_to_bf16_bits(v0), h1 = f32_to_bf16_bits(v1);
      *reinterpret_cast<uint32_t*>(out + off + db * 8) = (uint32_t)h0 | ((uint32_t)h1 << 16);
      if (out_lo) {
        const uint16_t l0 = f32_to_bf16_bits(v0 - __uint_as_float((uint32_t)h0 << 16));
        const uint16_t l1 = f32_to_bf16_bits(v1 - __uint_as_float((uint32_t)h1 << 16));
        *reinterpret_cast<uint32_t*>(out_lo + off + db * 8) = (uint32_t)l0 | ((uint32_t)l1 << 16);
      }
    }
  }
}

template <int D, int G, bool MID>
cudaError_t launch_t(const float* q, int T, int p0, int npad, const uint16_t* ws, int q_heads, int kv_heads,
                     uint16_t* out, uint16_t* out_lo, cudaStream_t s) {
  constexpr int S = D + 8;
  const size_t smem = ((size_t)2 * 16 * NWARP * S + (size_t)2 * (MID ? 4 : 2) * KT * S) * sizeof(uint16_t);
  auto kern = prefill_attn_mma_kernel<D, G, MID>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int rows_per_block = 16 * NWARP / G;
  dim3 grid((T + rows_per_block - 1) / rows_per_block, kv_heads);
  kern<<<grid, NWARP * 32, smem, s>>>(q, T, p0, npad, ws, q_heads, kv_heads, out, out_lo);
  return cudaGetLastError();
}

template <int D, bool MID>
cudaError_t launch_g(int G, const float* q, int T, int p0, int npad, const uint16_t* ws, int q_heads, int kv_heads,
                     uint16_t* out, uint16_t* out_lo, cudaStream_t s) {
  switch (G) {
    case 1: return launch_t<D, 1, MID>(q, T, p0, npad, ws, q_heads, kv_heads, out, out_lo, s);
    case 2: return launch_t<D, 2, MID>(q, T, p0, npad, ws, q_heads, kv_heads, out, out_lo, s);
    case 4: return launch_t<D, 4, MID>(q, T, p0, npad, ws, q_heads, kv_heads, out, out_lo, s);
    case 8: return launch_t<D, 8, MID>(q, T, p0, npad, ws, q_heads, kv_heads, out, out_lo, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

bool prefill_attn_mma_supported(int head_dim, int G) {
  return (head_dim == 64 || head_dim == 128) && (G == 1 || G == 2 || G == 4 || G == 8);
}

size_t prefill_attn_mma_workspace_bytes(int max_positions, int kv_heads, int head_dim) {
  const size_t npad = ((size_t)max_positions + KT - 1) / KT * KT;
  return (size_t)4 * kv_heads * npad * head_dim * sizeof(uint16_t);
}

cudaError_t launch_prefill_attention_mma(const float* q, int T, int p0, int slot, const int* block_table,
                                         int bt_stride, const void* kpool, const void* vpool, int kv_fp32,
                                         int page_size, int q_heads, int kv_heads, int head_dim, uint16_t* out_hi,
                                         uint16_t* out_lo, uint16_t* workspace, cudaStream_t s) {
  const int G = q_heads / kv_heads;
  if (!prefill_attn_mma_supported(head_dim, G) || !workspace) return cudaErrorInvalidValue;
  const int npos = p0 + T, npad = (npos + KT - 1) / KT * KT;
  dim3 sgrid(npad, kv_heads);
  if (kv_fp32)
    kv_split_kernel<float><<<sgrid, head_dim / 4, 0, s>>>(reinterpret_cast<const float*>(kpool),
                                                          reinterpret_cast<const float*>(vpool), slot, block_table,
                                                          bt_stride, page_size, kv_heads, head_dim, npos, npad, workspace);
  else
    kv_split_kernel<uint16_t><<<sgrid, head_dim / 4, 0, s>>>(reinterpret_cast<const uint16_t*>(kpool),
                                                             reinterpret_cast<const uint16_t*>(vpool), slot, block_table,
                                                             bt_stride, page_size, kv_heads, head_dim, npos, npad, workspace);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (head_dim == 64)
    return kv_fp32 ? launch_g<64, true>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s)
                   : launch_g<64, false>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s);
  return kv_fp32 ? launch_g<128, true>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s)
                 : launch_g<128, false>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s);
}

}  // namespace llmc
