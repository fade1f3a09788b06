// attn_decode.cu -- single-query (decode) GQA attention over the paged KV cache (bf16 or fp32).
//
//   out[h] = softmax(q[h] . K[0..ctx) / sqrt(d)) V[0..ctx)        fp32 softmax, online
//
// HBM/L2-bound on the KV read (ctx * kv_tok bytes per sequence per step),
// but at batch 1 the data is tiny (1.3 MB per layer for the 1B shape), so the kernel is built
// around its LATENCY chain rather than bandwidth:
//   * grid (nsplit, kv_heads, B), 16 warps per block.  Each warp owns whole 32-position tiles
//     (tile t of the block goes to warp t % 16) and streams them straight from the pages into
//     registers with fully coalesced 16-byte loads (LPR consecutive lanes cover one row, see
//     LaneMap); the same mapping serves the q.k dots and the PV product, one K/V read serves
//     every q head of the GQA group, and the softmax (lane = position) exchanges scores and p
//     values with it through a small per-warp shared-memory buffer.
//   * the 16 per-warp partial softmax states are merged through shared memory; with nsplit = 1
//     (contexts up to 2048) there is no global scratch, no atomic and no second pass.
//   * for longer contexts the block-level partials go to a scratch buffer and the last block of a
//     (sequence, kv head) -- one atomic counter, no atomics on data -- merges them.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the
// boundary; the providers behind it are remote HTTP calls).
#include "common.cuh"
#include "kernels.h"
#include <cstdlib>

#include "attn_tile.cuh"

namespace llmc {

namespace {

// warps per block: 16, or 8 when the per-head softmax state alone needs too many registers
template <int D, int G>
struct AttnCfg {
  static constexpr int AWARPS = (G * (D / 32 + 4) > 40) ? 8 : 16;
};

template <typename KVT>
struct KvTraits;
template <>
struct KvTraits<uint16_t> {
  static constexpr int EPV = 8;  // elements per 16-byte vector
  __device__ static __forceinline__ void unpack(const uint4& w, float (&f)[8]) {
    f[0] = bf16lo(w.x); f[1] = bf16hi(w.x); f[2] = bf16lo(w.y); f[3] = bf16hi(w.y);
    f[4] = bf16lo(w.z); f[5] = bf16hi(w.z); f[6] = bf16lo(w.w); f[7] = bf16hi(w.w);
  }
};
template <>
struct KvTraits<float> {
  static constexpr int EPV = 4;
  __device__ static __forceinline__ void unpack(const uint4& w, float (&f)[4]) {
    f[0] = __uint_as_float(w.x); f[1] = __uint_as_float(w.y); f[2] = __uint_as_float(w.z); f[3] = __uint_as_float(w.w);
  }
};

// AWARPS: warps per block.  Few (sequence, kv head) pairs: 16 (8) warps share one pair so that a long context is
// covered by many tiles in flight.  Many pairs (wide batches): 4 warps per block, so that several blocks are resident
// per SM and the staging / merge phases of one overlap the tile loads of the others.
// result element -> fp32 vector, or (wide decode path) directly the bf16 hi+lo operand pair of the O-projection GEMM
__device__ __forceinline__ void attn_store(const AttnDecodeArgs& a, size_t idx, float v) {
  if (a.out_hi) {
    const uint16_t h = f32_to_bf16_bits(v);
    a.out_hi[idx] = h;
    if (a.out_lo) a.out_lo[idx] = f32_to_bf16_bits(v - __uint_as_float((uint32_t)h << 16));
  } else {
    a.out[idx] = v;
  }
}

template <int D, int G, typename KVT, int AWARPS>
__global__ void __launch_bounds__(AWARPS * 32, AWARPS == 4 ? 4 : 1) attn_decode_kernel(const AttnDecodeArgs a) {
  using M = LaneMap<D, G, KVT>;
  constexpr int EPV = M::EPV, LPR = M::LPR, CPL = M::CPL, RPP = M::RPP, NP = M::NP, EPL = M::EPL, PPG = M::PPG;
  constexpr int DPL = D / 32;                       // dims per lane in the final merge
  constexpr int PS = D + 2;                         // partial record: o[D], m, l
  extern __shared__ __align__(16) float sm[];
  float* qs = sm;                                   // [G][D]
  float* part = sm + G * D;                         // [AWARPS][G][PS]
  float* pbuf = part + AWARPS * G * PS;             // [AWARPS][G][TP + 1]: scores / p values / corr per warp
  __shared__ bool s_last;

  pdl_launch_dependents();
  pdl_wait();

  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int sub = lane / LPR, c = lane % LPR;       // row within a pass, first chunk of this lane
  const int slot = a.slots[b];
  const int ctx = a.seq_len[slot] + a.ctx_extra;
  const int ntiles = (ctx + TP - 1) / TP;
  const int per = (ntiles + a.nsplit - 1) / a.nsplit;
  const int t0 = split * per, t1 = min(ntiles, t0 + per);
  const float scale = rsqrtf((float)D);
  const int QD = a.q_heads * D;
  const KVT* kpool = reinterpret_cast<const KVT*>(a.kpool);
  const KVT* vpool = reinterpret_cast<const KVT*>(a.vpool);
  const int* bt = a.block_table + (size_t)slot * a.bt_stride;
  float* pw = pbuf + (size_t)warp * G * (TP + 1);

  if (a.qkv_raw) {
    // Wide decode path: the QKV GEMM's raw output [B][(q_heads + 2 kv_heads) * D] comes straight here.  RoPE at this
    // sequence's position is applied to q while it is staged, and the block that owns the last tile rotates the new
    // K row and appends it and the V row to the sequence's page before any warp reads that tile (same arithmetic
    // as rope_append_batch_kernel, which this replaces).
    constexpr int hd2 = D / 2;
    const int pos = ctx - 1;
    const float* raw = a.qkv_raw + (size_t)b * (a.q_heads + 2 * a.kv_heads) * D;
    const float* rc = a.rope_cos + (size_t)pos * hd2;
    const float* rs = a.rope_sin + (size_t)pos * hd2;
    for (int i = tid; i < G * hd2; i += AWARPS * 32) {
      const int h = i / hd2, j = i - h * hd2;
      const float* src = raw + (size_t)(kvh * G + h) * D;
      const float v0 = src[j], v1 = src[j + hd2], cs = rc[j], sn = rs[j];
      qs[h * D + j] = v0 * cs - v1 * sn;
      qs[h * D + j + hd2] = v1 * cs + v0 * sn;
    }
    if ((ntiles - 1) / per == split) {
      const int page = bt[pos / a.page_size];
      const size_t eo = (((size_t)page * a.kv_heads + kvh) * a.page_size + (pos % a.page_size)) * D;
      KVT* kd = const_cast<KVT*>(kpool) + eo;
      KVT* vd = const_cast<KVT*>(vpool) + eo;
      const float* ks = raw + (size_t)(a.q_heads + kvh) * D;
      const float* vs = raw + (size_t)(a.q_heads + a.kv_heads + kvh) * D;
      for (int j = tid; j < hd2; j += AWARPS * 32) {
        const float k0 = ks[j], k1 = ks[j + hd2], cs = rc[j], sn = rs[j];
        const float r0 = k0 * cs - k1 * sn, r1 = k1 * cs + k0 * sn;
        if constexpr (sizeof(KVT) == 4) {
          kd[j] = r0; kd[j + hd2] = r1;
          vd[j] = vs[j]; vd[j + hd2] = vs[j + hd2];
        } else {
          kd[j] = f32_to_bf16_bits(r0); kd[j + hd2] = f32_to_bf16_bits(r1);
          vd[j] = f32_to_bf16_bits(vs[j]); vd[j + hd2] = f32_to_bf16_bits(vs[j + hd2]);
        }
      }
    }
  } else {
    for (int i = tid; i < G * D; i += AWARPS * 32) qs[i] = a.q[(size_t)b * QD + (size_t)kvh * G * D + i];
  }
  __syncthreads();   // q staged; the appended K/V row is visible to this block's loads

  float m[G], l[G], o[G][EPL];                      // m, l warp-uniform; o: partial over this lane's rows
#pragma unroll
  for (int h = 0; h < G; ++h) {
    m[h] = -INFINITY;
    l[h] = 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) o[h][e] = 0.f;
  }

  for (int t = t0 + warp; t < t1; t += AWARPS) {
    const int p0 = t * TP;
    const int valid = min(TP, ctx - p0);
    const int page = bt[p0 / a.page_size];
    const size_t base = (((size_t)page * a.kv_heads + kvh) * a.page_size + (p0 % a.page_size)) * D;
    const KVT* kl = kpool + base + c * EPV;          // + row * D + j * LPR * EPV
    const KVT* vl = vpool + base + c * EPV;

    attn_tile_step<D, G, KVT>(qs, pw, kl, vl, valid, scale, m, l, o);
  }

  // ---- reduce o over the row groups, then merge the per-warp states through shared memory
  attn_write_record<D, G, KVT>(part + (size_t)warp * G * PS, m, l, o);
  __syncthreads();
  if (warp < G) {
    const int h = warp, head = kvh * G + h;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < AWARPS; ++w) M = fmaxf(M, part[((size_t)w * G + h) * PS + D]);
    float L = 0.f, O[DPL];
#pragma unroll
    for (int i = 0; i < DPL; ++i) O[i] = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int w = 0; w < AWARPS; ++w) {
        const float* pp = part + ((size_t)w * G + h) * PS;
        const float mw = pp[D];
        const float wt = (mw == -INFINITY) ? 0.f : __expf(mw - M);
        L = fmaf(pp[D + 1], wt, L);
#pragma unroll
        for (int i = 0; i < DPL; ++i) O[i] = fmaf(pp[lane * DPL + i], wt, O[i]);
      }
    }
    if (a.nsplit == 1) {
      const float inv = 1.0f / L;
#pragma unroll
      for (int i = 0; i < DPL; ++i) attn_store(a, (size_t)b * QD + (size_t)head * D + lane * DPL + i, O[i] * inv);
    } else {
      float* gp = a.scratch + (((size_t)b * a.q_heads + head) * a.nsplit + split) * PS;
#pragma unroll
      for (int i = 0; i < DPL; ++i) gp[lane * DPL + i] = O[i];
      if (lane == 0) {
        gp[D] = M;
        gp[D + 1] = L;
      }
    }
  }
  if (a.nsplit == 1) return;

  // ---- long contexts: the last block of this (sequence, kv head) merges the block partials
  __threadfence();
  __syncthreads();
  unsigned* ctr = a.counters + (size_t)b * a.kv_heads + kvh;
  if (tid == 0) s_last = (atomicAdd(ctr, 1u) == (unsigned)(a.nsplit - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int e = tid; e < G * D; e += AWARPS * 32) {
    const int h = e / D, i = e - h * D;
    const int head = kvh * G + h;
    const float* gp = a.scratch + ((size_t)b * a.q_heads + head) * a.nsplit * PS;
    float M = -INFINITY;
    for (int sp = 0; sp < a.nsplit; ++sp) M = fmaxf(M, __ldcg(gp + (size_t)sp * PS + D));
    float L = 0.f, O = 0.f;
    for (int sp = 0; sp < a.nsplit; ++sp) {
      const float ms = __ldcg(gp + (size_t)sp * PS + D);
      const float wt = (ms == -INFINITY) ? 0.f : __expf(ms - M);
      L = fmaf(__ldcg(gp + (size_t)sp * PS + D + 1), wt, L);
      O = fmaf(__ldcg(gp + (size_t)sp * PS + i), wt, O);
    }
    attn_store(a, (size_t)b * QD + (size_t)head * D + i, O / L);
  }
  if (tid == 0) *ctr = 0u;
}

template <int D, int G, typename KVT, int AWARPS>
cudaError_t launch_w(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  const size_t smem = ((size_t)G * D + (size_t)AWARPS * G * (D + 2) + (size_t)AWARPS * G * (TP + 1)) * sizeof(float);
  auto kern = attn_decode_kernel<D, G, KVT, AWARPS>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a.nsplit, a.kv_heads, a.B);
  cfg.blockDim = dim3(AWARPS * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, a);
}

template <int D, int G, typename KVT>
cudaError_t launch_t(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  // two waves of blocks or more: small blocks, several resident per SM (measured: worse below that)
  if ((long long)a.nsplit * a.kv_heads * a.B >= 2 * 148) return launch_w<D, G, KVT, 4>(a, pdl, s);
  // In between (batches of about 4..36 sequences): 8 warps.  A 16-warp block at 128 registers per thread fills the
  // SM's register file, so the blocks of the NEXT kernel of the programmatic-launch chain (the O projection, which
  // streams its first weight tiles before it waits) cannot become resident beside it; with 8 warps they can.  Measured
  // 2-4 % per step at batches 8..32 on both shapes, contexts 512 and 2 048.  LLMC_ATTN_WARPS=16 restores 16.
  static const int aw = getenv("LLMC_ATTN_WARPS") ? atoi(getenv("LLMC_ATTN_WARPS")) : 8;
  if (aw == 8 && (long long)a.nsplit * a.kv_heads * a.B >= 32) return launch_w<D, G, KVT, 8>(a, pdl, s);
  return launch_w<D, G, KVT, AttnCfg<D, G>::AWARPS>(a, pdl, s);
}

template <int D, typename KVT>
cudaError_t launch_g(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  switch (a.q_heads / a.kv_heads) {
    case 1: return launch_t<D, 1, KVT>(a, pdl, s);
    case 2: return launch_t<D, 2, KVT>(a, pdl, s);
    case 4: return launch_t<D, 4, KVT>(a, pdl, s);
    case 8: return launch_t<D, 8, KVT>(a, pdl, s);
  }
  return cudaErrorInvalidValue;
}

template <typename KVT>
cudaError_t launch_d(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  switch (a.head_dim) {
    case 32: return launch_g<32, KVT>(a, pdl, s);
    case 64: return launch_g<64, KVT>(a, pdl, s);
    case 128: return launch_g<128, KVT>(a, pdl, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

size_t attn_decode_scratch_floats(int B, int q_heads, int nsplit, int head_dim) {
  return (size_t)B * q_heads * nsplit * (head_dim + 2);
}

// One block of 16 warps covers up to 64 tiles (2048 positions, 4 rounds per warp) before a
// second block per (sequence, kv head) pays for its global merge.
int attn_decode_nsplit(int max_ctx_tokens) {
  const int ntiles = (max_ctx_tokens + TP - 1) / TP;
  int n = (ntiles + 63) / 64;
  return n < 1 ? 1 : (n > ATTN_MAX_SPLIT ? ATTN_MAX_SPLIT : n);
}

cudaError_t launch_attn_decode(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  if (a.q_heads % a.kv_heads != 0 || a.page_size % TP != 0 || a.nsplit < 1 || a.nsplit > ATTN_MAX_SPLIT)
    return cudaErrorInvalidValue;
  return a.kv_fp32 ? launch_d<float>(a, pdl, s) : launch_d<uint16_t>(a, pdl, s);
}

}  // namespace llmc
