// prefill_attn_tc.cu -- causal GQA attention for the prefill on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Tensor-pipe bound (4 * n_h * d * T^2 / 2 flops per layer, causal).  Flash-attention
// structure, one CTA per (kv head, 128 / G query positions): the 128 accumulator rows are the G query heads of the
// GQA group x 128 / G positions, so one K/V tile serves the whole group.  Per 64-position K/V tile:
//     S  = Q K^T      tcgen05.mma, M = 128, N = 64,  operands in 128B-swizzled shared memory, S in TMEM (double buffered)
//     P  = softmax    four warps, one thread per row: tcgen05.ld of the row, running max / sum in registers (no
//                     shuffles), P written back to shared memory in the operand layout
//     PV = P V        tcgen05.mma, M = 128, N = D, into a third TMEM region; the row's thread folds it into its
//                     register accumulator with the softmax correction (no read-modify-write of TMEM)
// Warp roles: warp 0 = TMA producer (K and V^T tiles of the dense per-layer workspace, 2-stage mbarrier ring), warp 1 =
// TMEM allocation + the single MMA-issuing thread, warps 2-5 = softmax / epilogue.  S of tile j + 1 is issued before
// the wait for P of tile j, so the tensor pipe works on it while the softmax warps are busy with tile j.
//
// Precision is that of prefill_attn_mma.cu: every operand is a bf16 pair hi + mid and a
// product is three MMA chains into the same accumulator (hi*hi + mid*hi + hi*mid; two with a bf16 KV cache).
// kv_split_t_kernel writes the workspace: K planes [kv_head][position][d] (K-major operand of S) and V planes
// TRANSPOSED [kv_head][d][position] (K-major operand of PV), zero-padded to a multiple of 64 positions.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include "common.cuh"
#include "kernels.h"
#include "tc_common.cuh"

namespace llmc {

namespace {
using namespace tc;

constexpr int AKT = 64;        // positions per K/V tile
constexpr int AM = 128;        // accumulator rows per CTA
constexpr int ANST = 2;        // K/V stages
constexpr int ATHREADS = 192;

__device__ __forceinline__ uint32_t pk2(float lo, float hi) {   // lo -> bits 0..15
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& h, uint32_t& m) {
  h = pk2(x0, x1);
  m = pk2(x0 - __uint_as_float(h << 16), x1 - __uint_as_float(h & 0xffff0000u));
}
__device__ __forceinline__ void mbar_arrive1(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// byte offset of 16-byte chunk `c` of row `r` inside a [rows][64] bf16 tile in the 128B-swizzled K-major layout
// (what TMA writes with SWIZZLE_128B and make_smem_desc describes): 8-row groups of 1024 B, chunk index ^ row-in-group
__device__ __forceinline__ uint32_t sw128(int r, int c) {
  return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}

// ---- K/V pages of one slot -> dense split workspace.  Planes: khi, kmid [kv_heads][npad][D]; vhiT, vmidT [kv_heads][D][npad].
// grid (npad / 64, kv_heads), 256 threads: a block moves 64 positions of one head; V goes through shared memory so
// that both the page reads and the transposed writes are coalesced.
template <typename KVT, int D>
__global__ void __launch_bounds__(256) kv_split_t_kernel(const KVT* __restrict__ kpool, const KVT* __restrict__ vpool,
                                                         int slot, const int* __restrict__ block_table, int bt_stride,
                                                         int page_size, int kv_heads, int npos, int npad,
                                                         uint16_t* __restrict__ ws) {
  __shared__ float vs[64][D + 1];
  const int pb = blockIdx.x * 64, h = blockIdx.y, tid = threadIdx.x;
  const size_t plane = (size_t)kv_heads * npad * D;
  uint16_t* khi = ws;
  uint16_t* kmid = ws + plane;
  uint16_t* vhiT = ws + 2 * plane;
  uint16_t* vmidT = ws + 3 * plane;
  constexpr bool MID = sizeof(KVT) == 4;
  for (int e = tid; e < 64 * (D / 4); e += 256) {
    const int r = e / (D / 4), c = (e - r * (D / 4)) * 4;
    const int pos = pb + r;
    float4 k = make_float4(0.f, 0.f, 0.f, 0.f), v = k;
    if (pos < npos) {
      const int page = block_table[(size_t)slot * bt_stride + pos / page_size];
      const size_t src = (((size_t)page * kv_heads + h) * page_size + (pos % page_size)) * D + c;
      if constexpr (MID) {
        k = *reinterpret_cast<const float4*>(kpool + src);
        v = *reinterpret_cast<const float4*>(vpool + src);
      } else {
        const uint2 kb = *reinterpret_cast<const uint2*>(kpool + src), vb = *reinterpret_cast<const uint2*>(vpool + src);
        k = make_float4(bf16lo(kb.x), bf16hi(kb.x), bf16lo(kb.y), bf16hi(kb.y));
        v = make_float4(bf16lo(vb.x), bf16hi(vb.x), bf16lo(vb.y), bf16hi(vb.y));
      }
    }
    uint2 hh, mm;
    split2(k.x, k.y, hh.x, mm.x);
    split2(k.z, k.w, hh.y, mm.y);
    const size_t dst = ((size_t)h * npad + pos) * D + c;
    *reinterpret_cast<uint2*>(khi + dst) = hh;
    if (MID) *reinterpret_cast<uint2*>(kmid + dst) = mm;
    vs[r][c] = v.x; vs[r][c + 1] = v.y; vs[r][c + 2] = v.z; vs[r][c + 3] = v.w;
  }
  __syncthreads();
  // transposed write: row d of V^T, positions pb .. pb + 63 (128 contiguous bytes)
  for (int e = tid; e < D * 16; e += 256) {
    const int d = e >> 4, p4 = (e & 15) * 4;
    uint2 hh, mm;
    split2(vs[p4][d], vs[p4 + 1][d], hh.x, mm.x);
    split2(vs[p4 + 2][d], vs[p4 + 3][d], hh.y, mm.y);
    const size_t dst = ((size_t)h * D + d) * npad + pb + p4;
    *reinterpret_cast<uint2*>(vhiT + dst) = hh;
    if (MID) *reinterpret_cast<uint2*>(vmidT + dst) = mm;
  }
}

template <int D, bool MID>
struct AttnCfg {
  static constexpr int NPL = MID ? 2 : 1;                      // hi (+ mid) planes of K and of V
  static constexpr int Q_BYTES = AM * D * 2;                    // one Q plane: D / 64 blocks of [128][64]
  static constexpr int P_BYTES = AM * AKT * 2;                  // one P plane
  static constexpr int K_BYTES = AKT * D * 2;                   // one K plane of a stage: D / 64 blocks of [64][64]
  static constexpr int V_BYTES = D * AKT * 2;                   // one V^T plane of a stage: [D][64]
  static constexpr int STAGE_BYTES = NPL * (K_BYTES + V_BYTES);
  static constexpr int SMEM = 2 * Q_BYTES + 2 * P_BYTES + ANST * STAGE_BYTES + 1024 /*alignment*/ + 256 /*barriers*/;
  static constexpr int TMEM_COLS = 256;                         // S0 [0,64), S1 [64,128), PV [128, 128 + D)
};

template <int D, int G, bool MID>
__global__ void __launch_bounds__(ATHREADS, 1)
    prefill_attn_tc_kernel(const __grid_constant__ CUtensorMap tmKhi, const __grid_constant__ CUtensorMap tmKmid,
                           const __grid_constant__ CUtensorMap tmVhi, const __grid_constant__ CUtensorMap tmVmid,
                           const float* __restrict__ q, int T, int p0, int npad, int q_heads, int kv_heads,
                           uint16_t* __restrict__ out, uint16_t* __restrict__ out_lo) {
  using C = AttnCfg<D, MID>;
  constexpr int R = AM / G;                                     // query positions per CTA
  constexpr int DB = D / 64;                                    // 64-column blocks of a D-wide operand
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* base_ptr = smem_raw + (base - raw);
  const uint32_t q_hi = base, q_mid = base + C::Q_BYTES;
  const uint32_t p_hi = base + 2 * C::Q_BYTES, p_mid = p_hi + C::P_BYTES;
  const uint32_t st0 = p_mid + C::P_BYTES;
  const uint32_t bar_base = st0 + ANST * C::STAGE_BYTES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base_ptr + (bar_base - base) + 128);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (ANST + s); };
  auto s_full = [&](int b) { return bar_base + 8u * (2 * ANST + b); };
  const uint32_t p_full = bar_base + 8u * (2 * ANST + 2), pv_full = p_full + 8u, q_ready = p_full + 16u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kvh = blockIdx.y;
  const int rb = gridDim.x - 1 - blockIdx.x;                    // long (late) row blocks are scheduled first
  const int last_q = min(T - 1, rb * R + R - 1);
  const int ntiles = (p0 + last_q) / AKT + 1;

  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmKhi) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmVhi) : "memory");
    for (int s = 0; s < ANST; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(s_full(0), 1);
    mbar_init(s_full(1), 1);
    mbar_init(p_full, 128);
    mbar_init(pv_full, 1);
    mbar_init(q_ready, 128);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      for (int j = 0; j < ntiles; ++j) {
        const int s = j % ANST;
        const uint32_t ph = (j / ANST) & 1;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_expect_tx(full_bar(s), C::STAGE_BYTES);
        const uint32_t sk = st0 + s * C::STAGE_BYTES;            // K hi | K mid | V^T hi | V^T mid
        const uint32_t sv = sk + C::NPL * C::K_BYTES;
#pragma unroll
        for (int b = 0; b < DB; ++b) {
          tma_load_2d(sk + b * (AKT * 128), &tmKhi, b * 64, kvh * npad + j * AKT, full_bar(s));
          if (MID) tma_load_2d(sk + C::K_BYTES + b * (AKT * 128), &tmKmid, b * 64, kvh * npad + j * AKT, full_bar(s));
        }
        tma_load_2d(sv, &tmVhi, j * AKT, kvh * D, full_bar(s));
        if (MID) tma_load_2d(sv + C::V_BYTES, &tmVmid, j * AKT, kvh * D, full_bar(s));
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc(AM, AKT), idesc_o = make_idesc(AM, D);
      auto issue_s = [&](int j) {   // S_j = Q K_j^T into S buffer j & 1
        const int s = j % ANST;
        mbar_wait(full_bar(s), (uint32_t)((j / ANST) & 1));
        tc_fence_after();
        const uint32_t sk = st0 + s * C::STAGE_BYTES;
        const uint32_t d = tmem_base + (uint32_t)((j & 1) * AKT);
        uint32_t acc = 0u;
#pragma unroll
        for (int chain = 0; chain < (MID ? 3 : 2); ++chain) {
          // hi*hi, mid*hi, hi*mid
          const uint32_t qa = chain == 1 ? q_mid : q_hi;
          const uint32_t kb = sk + (chain == 2 ? C::K_BYTES : 0);
#pragma unroll
          for (int b = 0; b < DB; ++b)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              tc_mma_bf16(d, make_smem_desc(qa + b * (AM * 128) + k * 32), make_smem_desc(kb + b * (AKT * 128) + k * 32),
                          idesc_s, acc);
              acc = 1u;
            }
        }
        tc_commit(s_full(j & 1));
      };
      mbar_wait(q_ready, 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < ntiles; ++j) {
        if (j + 1 < ntiles) issue_s(j + 1);          // runs while the softmax warps work on tile j
        mbar_wait(p_full, (uint32_t)(j & 1));
        tc_fence_after();
        const int s = j % ANST;
        const uint32_t sv = st0 + s * C::STAGE_BYTES + C::NPL * C::K_BYTES;
        const uint32_t d = tmem_base + 2u * AKT;
        uint32_t acc = 0u;
#pragma unroll
        for (int chain = 0; chain < (MID ? 3 : 2); ++chain) {
          const uint32_t pa = chain == 1 ? p_mid : p_hi;
          const uint32_t vb = sv + (chain == 2 ? C::V_BYTES : 0);
#pragma unroll
          for (int k = 0; k < AKT / 16; ++k) {
            tc_mma_bf16(d, make_smem_desc(pa + k * 32), make_smem_desc(vb + k * 32), idesc_o, acc);
            acc = 1u;
          }
        }
        tc_commit(pv_full);
        tc_commit(empty_bar(s));   // K_j and V_j are no longer needed once S_j and PV_j have run
      }
    }
  } else {
    // ---------------------------------------------------------------- softmax / epilogue: one thread per row
    const int quad = warp & 3;                       // the TMEM lane quadrant this warp may access
    const int r = quad * 32 + lane;                  // accumulator row = TMEM lane
    const int h = r / R, t = r - h * R;
    const int qpos = rb * R + t;                     // query position within this call's T rows
    const bool live = qpos < T;
    const int lim = p0 + min(qpos, T - 1);           // last visible cache position
    const int head = kvh * G + h;
    const int QD = q_heads * D;
    const float scale = rsqrtf((float)D);
    const uint32_t tq = tmem_base + ((uint32_t)(quad * 32) << 16);
    // ---- Q row: scaled, split, written into the operand layout
    {
      const float* qr = q + (size_t)min(qpos, T - 1) * QD + (size_t)head * D;
#pragma unroll
      for (int c = 0; c < D / 8; ++c) {
        const float4 a = *reinterpret_cast<const float4*>(qr + c * 8);
        const float4 b = *reinterpret_cast<const float4*>(qr + c * 8 + 4);
        uint32_t hh[4], mm[4];
        split2(a.x * scale, a.y * scale, hh[0], mm[0]);
        split2(a.z * scale, a.w * scale, hh[1], mm[1]);
        split2(b.x * scale, b.y * scale, hh[2], mm[2]);
        split2(b.z * scale, b.w * scale, hh[3], mm[3]);
        const uint32_t off = (uint32_t)((c >> 3) * (AM * 128)) + sw128(r, c & 7);
        sts128(q_hi + off, hh[0], hh[1], hh[2], hh[3]);
        sts128(q_mid + off, mm[0], mm[1], mm[2], mm[3]);
      }
      fence_async_smem();
      mbar_arrive1(q_ready);
    }
    float o[D];
#pragma unroll
    for (int i = 0; i < D; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < ntiles; ++j) {
      mbar_wait(s_full(j & 1), (uint32_t)((j >> 1) & 1));
      tc_fence_after();
      uint32_t sv[AKT];
      {
        uint32_t a[32], b[32];
        tc_ld32(tq + (uint32_t)((j & 1) * AKT), a);
        tc_ld32(tq + (uint32_t)((j & 1) * AKT + 32), b);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) { sv[i] = a[i]; sv[32 + i] = b[i]; }
      }
      const int pbase = j * AKT;
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < AKT; ++i) {
        float s = __uint_as_float(sv[i]);
        if (pbase + i > lim) s = -INFINITY;
        sv[i] = __float_as_uint(s);
        mx = fmaxf(mx, s);
      }
      const float mn = fmaxf(m, mx);              // position 0 is visible to every row, so mn is finite
      const float corr = __expf(m - mn);
      m = mn;
      float rsum = 0.f;
#pragma unroll
      for (int c = 0; c < AKT / 8; ++c) {
        float p[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          p[i] = __expf(__uint_as_float(sv[c * 8 + i]) - m);
          rsum += p[i];
        }
        uint32_t hh[4], mm[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split2(p[2 * i], p[2 * i + 1], hh[i], mm[i]);
        const uint32_t off = sw128(r, c);
        sts128(p_hi + off, hh[0], hh[1], hh[2], hh[3]);
        sts128(p_mid + off, mm[0], mm[1], mm[2], mm[3]);
      }
      l = l * corr + rsum;
      tc_fence_before();
      fence_async_smem();
      mbar_arrive1(p_full);
      // ---- fold PV_j into the register accumulator
      mbar_wait(pv_full, (uint32_t)(j & 1));
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t v[32];
        tc_ld32(tq + 2u * AKT + (uint32_t)(c * 32), v);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], corr, __uint_as_float(v[i]));
      }
      tc_fence_before();
    }
    // ---- normalise and store as bf16 hi (+ lo)
    if (live) {
      const float inv = 1.0f / l;
      const size_t off = (size_t)qpos * QD + (size_t)head * D;
#pragma unroll
      for (int c = 0; c < D / 8; ++c) {
        uint32_t hh[4], ll[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split2(o[c * 8 + 2 * i] * inv, o[c * 8 + 2 * i + 1] * inv, hh[i], ll[i]);
        *reinterpret_cast<uint4*>(out + off + c * 8) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        if (out_lo) *reinterpret_cast<uint4*>(out_lo + off + c * 8) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
  }
}

// 2-D bf16 tensor [rows][cols], box [box_rows][64 cols], 128B swizzle (tc::make_map with an explicit box)
template <int D, int G, bool MID>
cudaError_t launch_t(const float* q, int T, int p0, int npad, const uint16_t* ws, int q_heads, int kv_heads,
                     uint16_t* out, uint16_t* out_lo, cudaStream_t s) {
  using C = AttnCfg<D, MID>;
  const size_t plane = (size_t)kv_heads * npad * D;
  CUtensorMap mkh, mkm, mvh, mvm;
  // K planes: rows = kv_heads * npad positions, cols = D; V^T planes: rows = kv_heads * D, cols = npad positions
  if (!make_map(&mkh, ws, (uint64_t)kv_heads * npad, (uint64_t)D, AKT) ||
      !make_map(&mkm, ws + plane, (uint64_t)kv_heads * npad, (uint64_t)D, AKT) ||
      !make_map(&mvh, ws + 2 * plane, (uint64_t)kv_heads * D, (uint64_t)npad, D) ||
      !make_map(&mvm, ws + 3 * plane, (uint64_t)kv_heads * D, (uint64_t)npad, D))
    return cudaErrorInvalidValue;
  auto kern = prefill_attn_tc_kernel<D, G, MID>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
  if (e != cudaSuccess) return e;
  constexpr int R = AM / G;
  dim3 grid((T + R - 1) / R, kv_heads);
  kern<<<grid, ATHREADS, C::SMEM, s>>>(mkh, mkm, mvh, mvm, q, T, p0, npad, q_heads, kv_heads, out, out_lo);
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

bool prefill_attn_tc_supported(int head_dim, int G) {
  return (head_dim == 64 || head_dim == 128) && (G == 1 || G == 2 || G == 4 || G == 8);
}

cudaError_t launch_prefill_attention_tc(const float* q, int T, int p0, int slot, const int* block_table,
                                        int bt_stride, const void* kpool, const void* vpool, int kv_fp32,
                                        int page_size, int q_heads, int kv_heads, int head_dim, uint16_t* out_hi,
                                        uint16_t* out_lo, uint16_t* workspace, cudaStream_t s) {
  const int G = q_heads / kv_heads;
  if (!prefill_attn_tc_supported(head_dim, G) || !workspace) return cudaErrorInvalidValue;
  const int npos = p0 + T;
  const int npad = (npos + AKT - 1) / AKT * AKT;
  dim3 sg(npad / 64, kv_heads);
  if (head_dim == 64) {
    if (kv_fp32)
      kv_split_t_kernel<float, 64><<<sg, 256, 0, s>>>(reinterpret_cast<const float*>(kpool), reinterpret_cast<const float*>(vpool),
                                                      slot, block_table, bt_stride, page_size, kv_heads, npos, npad, workspace);
    else
      kv_split_t_kernel<uint16_t, 64><<<sg, 256, 0, s>>>(reinterpret_cast<const uint16_t*>(kpool),
                                                         reinterpret_cast<const uint16_t*>(vpool), slot, block_table,
                                                         bt_stride, page_size, kv_heads, npos, npad, workspace);
  } else {
    if (kv_fp32)
      kv_split_t_kernel<float, 128><<<sg, 256, 0, s>>>(reinterpret_cast<const float*>(kpool), reinterpret_cast<const float*>(vpool),
                                                       slot, block_table, bt_stride, page_size, kv_heads, npos, npad, workspace);
    else
      kv_split_t_kernel<uint16_t, 128><<<sg, 256, 0, s>>>(reinterpret_cast<const uint16_t*>(kpool),
                                                          reinterpret_cast<const uint16_t*>(vpool), slot, block_table,
                                                          bt_stride, page_size, kv_heads, npos, npad, workspace);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (head_dim == 64)
    return kv_fp32 ? launch_g<64, true>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s)
                   : launch_g<64, false>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s);
  return kv_fp32 ? launch_g<128, true>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s)
                 : launch_g<128, false>(G, q, T, p0, npad, workspace, q_heads, kv_heads, out_hi, out_lo, s);
}

}  // namespace llmc
