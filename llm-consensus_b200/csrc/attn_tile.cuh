// attn_tile.cuh -- one 32-position tile of single-query GQA attention for one warp: coalesced K/V
// reads straight from the pages, online softmax, shared by the stand-alone decode attention kernel
// (attn_decode.cu) and the one-kernel decode step (decode_mega.cu).
#pragma once
#include "common.cuh"

namespace llmc {

constexpr int TP = 32;  // positions per tile

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


// Folds one tile into the warp's running softmax state.
//   qs  [G][D] fp32 q of the GQA group (shared memory)         pw  [G][TP+1] per-warp exchange buffer (smem)
//   kl / vl: the tile's first K / V row, already offset by this lane's first chunk (c * EPV)
//   valid: positions of the tile that exist; m, l warp-uniform; o partial over this lane's rows
//   kl / vl carry no __restrict__: a caller may have appended the newest row to the pools earlier in the same kernel
//   (attn_decode.cu, decode_flow.cu), and const + __restrict__ would let the compiler read through the
//   non-coherent path
template <bool CG>
__device__ __forceinline__ uint4 ld16(const void* p) {
  // CG: read through L2 only (the caller's kernel communicates this data between SMs)
  if constexpr (CG) return __ldcg(reinterpret_cast<const uint4*>(p));
  else return *reinterpret_cast<const uint4*>(p);
}

// LDF: 16-byte loads in flight per phase (a kernel with registers to spare raises it so that a whole K or V tile is
// requested at once: one memory round trip per phase instead of two)
template <int D, int G, typename KVT, bool CG = false, int LDF = LaneMap<D, G, KVT>::LD>
__device__ __forceinline__ void attn_tile_step(const float* __restrict__ qs, float* __restrict__ pw,
                                               const KVT* kl, const KVT* vl, int valid,
                                               float scale, float (&m)[G], float (&l)[G],
                                               float (&o)[G][LaneMap<D, G, KVT>::EPL]) {
  using M = LaneMap<D, G, KVT>;
  constexpr int EPV = M::EPV, LPR = M::LPR, CPL = M::CPL, RPP = M::RPP, NP = M::NP, EPL = M::EPL;
  constexpr int PPG = (LDF / CPL) < NP ? ((LDF / CPL) < 1 ? 1 : (LDF / CPL)) : NP;
  const int lane = threadIdx.x & 31;
  const int sub = lane / LPR, c = lane % LPR;
    // ---- scores: partial dots over this lane's chunks, reduced over the LPR lanes of a row
#pragma unroll
    for (int g0 = 0; g0 < NP; g0 += PPG) {
      uint4 kv[PPG][CPL];
#pragma unroll
      for (int pp = 0; pp < PPG; ++pp) {
        const int row = min((g0 + pp) * RPP + sub, valid - 1);   // rows past the end alias the last valid one
#pragma unroll
        for (int j = 0; j < CPL; ++j)
          kv[pp][j] = ld16<CG>(kl + (size_t)row * D + j * LPR * EPV);
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
          vv[pp][j] = ld16<CG>(vl + (size_t)row * D + j * LPR * EPV);
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

// ---- the same tile step split in two, for callers that can fetch a tile long before they may use it (the one-kernel
// decode step fetches each warp's first K/V tile while it waits for the grid barrier that publishes q): the whole tile
// (K and V) sits in registers, NP * CPL 16-byte chunks each.
template <int D, int G, typename KVT>
struct TileRegs {
  using M = LaneMap<D, G, KVT>;
  uint4 k[M::NP][M::CPL];
  uint4 v[M::NP][M::CPL];
};

// rows: positions of the tile that may be read now (rows at or past it alias row rows - 1, as in attn_tile_step)
template <int D, int G, typename KVT, bool CG>
__device__ __forceinline__ void attn_tile_fetch(const KVT* kl, const KVT* vl, int rows, TileRegs<D, G, KVT>& t) {
  using M = LaneMap<D, G, KVT>;
  const int sub = (threadIdx.x & 31) / M::LPR;
#pragma unroll
  for (int pp = 0; pp < M::NP; ++pp) {
    const int row = min(pp * M::RPP + sub, rows - 1);
#pragma unroll
    for (int j = 0; j < M::CPL; ++j) {
      t.k[pp][j] = ld16<CG>(kl + (size_t)row * D + j * M::LPR * M::EPV);
      t.v[pp][j] = ld16<CG>(vl + (size_t)row * D + j * M::LPR * M::EPV);
    }
  }
}
// (re)loads one row of a fetched tile: the lanes that hold row `r` read it again
template <int D, int G, typename KVT, bool CG>
__device__ __forceinline__ void attn_tile_fetch_row(const KVT* kl, const KVT* vl, int r, TileRegs<D, G, KVT>& t) {
  using M = LaneMap<D, G, KVT>;
  const int sub = (threadIdx.x & 31) / M::LPR;
#pragma unroll
  for (int pp = 0; pp < M::NP; ++pp) {
    if (pp * M::RPP + sub == r) {
#pragma unroll
      for (int j = 0; j < M::CPL; ++j) {
        t.k[pp][j] = ld16<CG>(kl + (size_t)r * D + j * M::LPR * M::EPV);
        t.v[pp][j] = ld16<CG>(vl + (size_t)r * D + j * M::LPR * M::EPV);
      }
    }
  }
}
// folds a fetched tile into the running softmax state: the arithmetic of attn_tile_step, operation for operation
template <int D, int G, typename KVT>
__device__ __forceinline__ void attn_tile_fold(const float* __restrict__ qs, float* __restrict__ pw,
                                               const TileRegs<D, G, KVT>& t, int valid, float scale, float (&m)[G],
                                               float (&l)[G], float (&o)[G][LaneMap<D, G, KVT>::EPL]) {
  using M = LaneMap<D, G, KVT>;
  constexpr int EPV = M::EPV, LPR = M::LPR, CPL = M::CPL, RPP = M::RPP, NP = M::NP, EPL = M::EPL;
  const int lane = threadIdx.x & 31;
  const int sub = lane / LPR, c = lane % LPR;
#pragma unroll
  for (int pp = 0; pp < NP; ++pp) {
    float sc[G];
#pragma unroll
    for (int h = 0; h < G; ++h) sc[h] = 0.f;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      float kf[EPV];
      unpack_chunk<KVT>(t.k[pp][j], kf);
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
      if (c == 0) pw[h * (TP + 1) + pp * RPP + sub] = sc[h] * scale;
    }
  }
  __syncwarp();
#pragma unroll
  for (int h = 0; h < G; ++h) {
    const float sh = (lane < valid) ? pw[h * (TP + 1) + lane] : -INFINITY;
    const float mn = fmaxf(m[h], warp_max(sh));
    const float corr = __expf(m[h] - mn);
    const float ph = (lane < valid) ? __expf(sh - mn) : 0.f;
    l[h] = l[h] * corr + warp_sum(ph);
    m[h] = mn;
    pw[h * (TP + 1) + lane] = ph;
#pragma unroll
    for (int e = 0; e < EPL; ++e) o[h][e] *= corr;
  }
  __syncwarp();
#pragma unroll
  for (int pp = 0; pp < NP; ++pp) {
    const int row = pp * RPP + sub;
    const bool live = row < valid;
    float pr[G];
#pragma unroll
    for (int h = 0; h < G; ++h) pr[h] = live ? pw[h * (TP + 1) + row] : 0.f;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      float vf[EPV];
      unpack_chunk<KVT>(t.v[pp][j], vf);
#pragma unroll
      for (int h = 0; h < G; ++h)
#pragma unroll
        for (int e = 0; e < EPV; ++e) o[h][j * EPV + e] = fmaf(pr[h], vf[e], o[h][j * EPV + e]);
    }
  }
  __syncwarp();
}

// Sums the per-lane partial outputs over the row groups of the lane map and writes the warp's state
// record [G][D + 2] = (o[D], m, l) per head.
template <int D, int G, typename KVT>
__device__ __forceinline__ void attn_write_record(float* __restrict__ rec, float (&m)[G], float (&l)[G],
                                                  float (&o)[G][LaneMap<D, G, KVT>::EPL]) {
  using M = LaneMap<D, G, KVT>;
  constexpr int EPV = M::EPV, LPR = M::LPR, CPL = M::CPL, EPL = M::EPL, PS = D + 2;
  const int lane = threadIdx.x & 31;
  const int sub = lane / LPR, c = lane % LPR;
#pragma unroll
  for (int h = 0; h < G; ++h)
#pragma unroll
    for (int e = 0; e < EPL; ++e)
#pragma unroll
      for (int off = LPR; off < 32; off <<= 1) o[h][e] += __shfl_xor_sync(0xffffffffu, o[h][e], off);
  if (sub == 0) {
#pragma unroll
    for (int h = 0; h < G; ++h) {
      float* pp = rec + (size_t)h * PS;
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
}

}  // namespace llmc
