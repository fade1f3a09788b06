// attn_common.cuh -- the per-warp tile step shared by the decode and prefill attention kernels.
#pragma once
#include "common.cuh"

namespace llmc {

constexpr int ATT_TP = 32;  // positions per K/V tile

// One tile step for one q head handled by one warp.  K/V tiles are in smem with padded rows.
//   qs  : this head's q in smem, fp32 [d]
//   lane = position within the tile for the scores; lane = dim slice for the PV product
template <int D>
__device__ __forceinline__ void attend_tile(const float* __restrict__ qs, const uint16_t* __restrict__ ks,
                                            const uint16_t* __restrict__ vs, int kstride, int valid,
                                            float scale, float& m, float& l, float (&o)[D / 32]) {
  const int lane = threadIdx.x & 31;
  float s = -INFINITY;
  if (lane < valid) {
    const uint4* kr = reinterpret_cast<const uint4*>(ks + (size_t)lane * kstride);
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      uint4 w = kr[i];
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
    const uint16_t* vr = vs + (size_t)t * kstride + lane * (D / 32);
    if (D == 64) {
      uint32_t w = *reinterpret_cast<const uint32_t*>(vr);
      o[0] = fmaf(pt, bf16lo(w), o[0]);
      o[1] = fmaf(pt, bf16hi(w), o[1]);
    } else if (D == 128) {
      uint2 w = *reinterpret_cast<const uint2*>(vr);
      o[0] = fmaf(pt, bf16lo(w.x), o[0]);
      o[1] = fmaf(pt, bf16hi(w.x), o[1]);
      o[2] = fmaf(pt, bf16lo(w.y), o[2]);
      o[3] = fmaf(pt, bf16hi(w.y), o[3]);
    } else {  // D == 32
      o[0] = fmaf(pt, __uint_as_float((uint32_t)vr[0] << 16), o[0]);
    }
  }
}


}  // namespace llmc
