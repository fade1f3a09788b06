// attn_common.cuh -- the per-warp tile step of the prefill attention kernel: K/V tile staged in
// shared memory (bf16 or fp32, padded rows), one q head per call, online softmax.
#pragma once
#include "common.cuh"

namespace llmc {

constexpr int ATT_TP = 32;  // positions per K/V tile

// padded row stride (elements) that makes lane-per-row 128-bit reads conflict-free
template <typename KVT, int D>
struct TileStride {
  static constexpr int value = D + 16 / (int)sizeof(KVT);
};

//   qs   : this head's q in smem, fp32 [D]
//   lane = position within the tile for the scores; lane = dim slice for the PV product
template <int D, typename KVT>
__device__ __forceinline__ void attend_tile(const float* __restrict__ qs, const KVT* __restrict__ ks,
                                            const KVT* __restrict__ vs, int valid, float scale, float& m, float& l,
                                            float (&o)[D / 32]) {
  constexpr int KS = TileStride<KVT, D>::value;
  constexpr int EPV = 16 / (int)sizeof(KVT);
  const int lane = threadIdx.x & 31;
  float s = -INFINITY;
  if (lane < valid) {
    const uint4* kr = reinterpret_cast<const uint4*>(ks + (size_t)lane * KS);
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < D / EPV; ++i) {
      const uint4 w = kr[i];
      if constexpr (sizeof(KVT) == 2) {
        const float4 a = *reinterpret_cast<const float4*>(qs + i * 8);
        const float4 b = *reinterpret_cast<const float4*>(qs + i * 8 + 4);
        acc = fmaf(bf16lo(w.x), a.x, acc);
        acc = fmaf(bf16hi(w.x), a.y, acc);
        acc = fmaf(bf16lo(w.y), a.z, acc);
        acc = fmaf(bf16hi(w.y), a.w, acc);
        acc = fmaf(bf16lo(w.z), b.x, acc);
        acc = fmaf(bf16hi(w.z), b.y, acc);
        acc = fmaf(bf16lo(w.w), b.z, acc);
        acc = fmaf(bf16hi(w.w), b.w, acc);
      } else {
        const float4 a = *reinterpret_cast<const float4*>(qs + i * 4);
        acc = fmaf(__uint_as_float(w.x), a.x, acc);
        acc = fmaf(__uint_as_float(w.y), a.y, acc);
        acc = fmaf(__uint_as_float(w.z), a.z, acc);
        acc = fmaf(__uint_as_float(w.w), a.w, acc);
      }
    }
    s = acc * scale;
  }
  const float mt = warp_max(s);
  const float mn = fmaxf(m, mt);          // valid >= 1 so mn is finite
  const float corr = __expf(m - mn);      // m = -inf on the first tile -> 0
  const float p = (lane < valid) ? __expf(s - mn) : 0.f;
  l = l * corr + warp_sum(p);
  m = mn;
#pragma unroll
  for (int j = 0; j < D / 32; ++j) o[j] *= corr;
  // PV: lane owns dims [lane*(D/32), +D/32)
  for (int t = 0; t < valid; ++t) {
    const float pt = __shfl_sync(0xffffffffu, p, t);
    const KVT* vr = vs + (size_t)t * KS + lane * (D / 32);
#pragma unroll
    for (int j = 0; j < D / 32; ++j) {
      float vv;
      if constexpr (sizeof(KVT) == 2) vv = __uint_as_float((uint32_t)vr[j] << 16);
      else vv = vr[j];
      o[j] = fmaf(pt, vv, o[j]);
    }
  }
}

}  // namespace llmc
