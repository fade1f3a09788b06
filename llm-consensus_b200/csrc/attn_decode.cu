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

namespace llmc {

namespace {

constexpr int TP = 32;        // positions per tile
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

// Lane mapping shared by the score (K) and the PV (V) phase: a K/V row is split into 16-byte
// chunks; LPR consecutive lanes read consecutive chunks of one row (fully coalesced, 128+ B per
// row per request) and the warp covers RPP = 32/LPR rows per pass.  Each lane therefore sees
// EPL elements of every (32/RPP)-th row: it accumulates partial q.k dots (reduced over the LPR
// lanes with shuffles) and partial outputs o[head][its dims] (reduced over the RPP row groups
// once, at the very end).
template <int D, int G, typename KVT>
struct LaneMap {
  static constexpr int EPV = 16 / (int)sizeof(KVT);          // elements per 16-byte chunk
  static constexpr int RC = D / EPV;                         // chunks per row
  static constexpr int WANT = (G * D / 32) > 8 ? (G * D / 32) : 8;
  static constexpr int LPR = RC < WANT ? RC : (WANT > 32 ? 32 : WANT);  // lanes per row
  static constexpr int CPL = RC / LPR;                       // chunks per lane per row
  static constexpr int RPP = 32 / LPR;                       // rows per pass
  static constexpr int NP = TP / RPP;                        // passes per tile
  static constexpr int EPL = CPL * EPV;                      // elements per lane per row
  static constexpr int LD = 8;                               // 16-byte loads in flight per phase
  static constexpr int PPG = (LD / CPL) < NP ? (LD / CPL) : NP;  // passes per register group
};

template <typename KVT>
__device__ __forceinline__ void unpack_chunk(const uint4& w, float (&f)[16 / sizeof(KVT)]) {
  if constexpr (sizeof(KVT) == 2) {
    f[0] = bf16lo(w.x); f[1] = bf16hi(w.x); f[2] = bf16lo(w.y); f[3] = bf16hi(w.y);
    f[4] = bf16lo(w.z); f[5] = bf16hi(w.z); f[6] = bf16lo(w.w); f[7] = bf16hi(w.w);
  } else {
    f[0] = __uint_as_float(w.x); f[1] = __uint_as_float(w.y); f[2] = __uint_as_float(w.z); f[3] = __uint_as_float(w.w);
  }
}

template <int D, int G, typename KVT>
__global__ void __launch_bounds__(AttnCfg<D, G>::AWARPS * 32) attn_decode_kernel(const AttnDecodeArgs a) {
  constexpr int AWARPS = AttnCfg<D, G>::AWARPS;
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

  for (int i = tid; i < G * D; i += AWARPS * 32) qs[i] = a.q[(size_t)b * QD + (size_t)kvh * G * D + i];
  __syncthreads();

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

    // ---- scores: partial dots over this lane's chunks, reduced over the LPR lanes of a row
#pragma unroll
    for (int g0 = 0; g0 < NP; g0 += PPG) {
      uint4 kv[PPG][CPL];
#pragma unroll
      for (int pp = 0; pp < PPG; ++pp) {
        const int row = min((g0 + pp) * RPP + sub, valid - 1);   // rows past the end alias the last valid one
#pragma unroll
        for (int j = 0; j < CPL; ++j)
          kv[pp][j] = *reinterpret_cast<const uint4*>(kl + (size_t)row * D + j * LPR * EPV);
      }
#pragma unroll
      for (int pp = 0; pp < PPG; ++pp) {
        float sc[G];
#pragma unroll
        for (int h = 0; h < G; ++h) sc[h] = 0.f;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
          float kf[EPV];
          unpack_chunk<KVT>(kv[pp][j], kf);
#pragma unroll
          for (int h = 0; h < G; ++h) {
            const float* qh = qs + h * D + (c + j * LPR) * EPV;
#pragma unroll
            for (int e = 0; e < EPV; e += 4) {
              const float4 qv = *reinterpret_cast<const float4*>(qh + e);
              sc[h] = fmaf(kf[e], qv.x, sc[h]);
              sc[h] = fmaf(kf[e + 1], qv.y, sc[h]);
              sc[h] = fmaf(kf[e + 2], qv.z, sc[h]);
              sc[h] = fmaf(kf[e + 3], qv.w, sc[h]);
            }
          }
        }
#pragma unroll
        for (int h = 0; h < G; ++h) {
#pragma unroll
          for (int off = LPR / 2; off > 0; off >>= 1) sc[h] += __shfl_xor_sync(0xffffffffu, sc[h], off);
          if (c == 0) pw[h * (TP + 1) + (g0 + pp) * RPP + sub] = sc[h] * scale;
        }
      }
    }
    __syncwarp();
    // ---- online softmax: lane = position
#pragma unroll
    for (int h = 0; h < G; ++h) {
      const float sh = (lane < valid) ? pw[h * (TP + 1) + lane] : -INFINITY;
      const float mn = fmaxf(m[h], warp_max(sh));      // valid >= 1, so finite
      const float corr = __expf(m[h] - mn);            // first tile: exp(-inf) = 0
      const float ph = (lane < valid) ? __expf(sh - mn) : 0.f;
      l[h] = l[h] * corr + warp_sum(ph);
      m[h] = mn;
      pw[h * (TP + 1) + lane] = ph;
#pragma unroll
      for (int e = 0; e < EPL; ++e) o[h][e] *= corr;
    }
    __syncwarp();
    // ---- PV: same lane mapping; o[h][.] += p[h][row] * V[row][this lane's chunks]
#pragma unroll
    for (int g0 = 0; g0 < NP; g0 += PPG) {
      uint4 vv[PPG][CPL];
#pragma unroll
      for (int pp = 0; pp < PPG; ++pp) {
        const int row = min((g0 + pp) * RPP + sub, valid - 1);
#pragma unroll
        for (int j = 0; j < CPL; ++j)
          vv[pp][j] = *reinterpret_cast<const uint4*>(vl + (size_t)row * D + j * LPR * EPV);
      }
#pragma unroll
      for (int pp = 0; pp < PPG; ++pp) {
        const int row = (g0 + pp) * RPP + sub;
        const bool live = row < valid;                 // aliased rows contribute nothing
        float pr[G];
#pragma unroll
        for (int h = 0; h < G; ++h) pr[h] = live ? pw[h * (TP + 1) + row] : 0.f;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
          float vf[EPV];
          unpack_chunk<KVT>(vv[pp][j], vf);
#pragma unroll
          for (int h = 0; h < G; ++h)
#pragma unroll
            for (int e = 0; e < EPV; ++e) o[h][j * EPV + e] = fmaf(pr[h], vf[e], o[h][j * EPV + e]);
        }
      }
    }
    __syncwarp();                                      // p values consumed before the next tile's scores
  }

  // ---- reduce o over the RPP row groups, then merge the per-warp states through shared memory
#pragma unroll
  for (int h = 0; h < G; ++h)
#pragma unroll
    for (int e = 0; e < EPL; ++e)
#pragma unroll
      for (int off = LPR; off < 32; off <<= 1) o[h][e] += __shfl_xor_sync(0xffffffffu, o[h][e], off);
  if (sub == 0) {
#pragma unroll
    for (int h = 0; h < G; ++h) {
      float* pp = part + ((size_t)warp * G + h) * PS;
#pragma unroll
      for (int j = 0; j < CPL; ++j)
#pragma unroll
        for (int e = 0; e < EPV; ++e) pp[(c + j * LPR) * EPV + e] = o[h][j * EPV + e];
      if (c == 0) {
        pp[D] = m[h];
        pp[D + 1] = l[h];
      }
    }
  }
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
      for (int i = 0; i < DPL; ++i) a.out[(size_t)b * QD + (size_t)head * D + lane * DPL + i] = O[i] * inv;
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
    a.out[(size_t)b * QD + (size_t)head * D + i] = O / L;
  }
  if (tid == 0) *ctr = 0u;
}

template <int D, int G, typename KVT>
cudaError_t launch_t(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  constexpr int AWARPS = AttnCfg<D, G>::AWARPS;
  const size_t smem = ((size_t)G * D + (size_t)AWARPS * G * (D + 2) + (size_t)AWARPS * G * (TP + 1)) * sizeof(float);
  auto kern = attn_decode_kernel<D, G, KVT>;
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
