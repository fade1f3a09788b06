// gemm_tcgen05.cu -- the prefill GEMM on the 5th-generation tensor cores (sm_100a):
//   C[T][N] = A[T][K] * W[N][K]^T,  bf16 operands, fp32 accumulation in TMEM.
//
// Tensor-pipe bound (2*T*N*K flops per call).  One CTA per 128 x BN output
// tile, warp-specialised:
//   warp 0      TMA producer: cp.async.bulk.tensor 2-D loads of a 128x64 A tile and a BNx64 W
//               tile per k-step into a 128B-swizzled shared-memory ring (mbarrier full/empty)
//   warp 1      allocates TMEM; one elected lane issues tcgen05.mma (M=128, N=BN, K=16 per
//               instruction, 4 per k-step) with shared-memory operand descriptors; commits
//               release the ring slots and finally signal the epilogue
//   warps 2-5   epilogue: tcgen05.ld the accumulator (each warp its own TMEM lane quadrant),
//               apply the fused epilogue (bf16/f32 store, residual add, SwiGLU) and store
// Both operands are K-major, so W is used exactly as the decode GEMV streams it.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the
// boundary; the providers behind it are remote HTTP calls).
#include "common.cuh"
#include "kernels.h"
#include "tc_common.cuh"

namespace llmc {

namespace {

constexpr int BM = 128, BK = 64, UMMA_K = 16;
constexpr int GEMM_THREADS = 192;

using namespace tc;

// SPLIT: A is given as a bf16 pair A_hi + A_lo (16 mantissa bits); each k-step then issues the
// MMAs twice against the same W tile, accumulating A_hi W + A_lo W in TMEM.
template <int BN, bool SPLIT>
struct TileCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int A_TILES = SPLIT ? 2 : 1;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_TILES * A_BYTES + B_BYTES;
  static constexpr int STAGES = (200 * 1024 / STAGE_BYTES) > 6 ? 6 : (200 * 1024 / STAGE_BYTES);
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;
};

template <int BN, int EPI, bool SPLIT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
    gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmAlo,
                   const __grid_constant__ CUtensorMap tmW, void* __restrict__ out, void* __restrict__ out_lo, int T,
                   int N, int K) {
  using C = TileCfg<BN, SPLIT>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // SWIZZLE_128B operands need 1024-byte alignment
  unsigned char* base_ptr = smem_raw + (base - raw);
  const uint32_t bar_base = base + C::STAGES * C::STAGE_BYTES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base_ptr + C::STAGES * C::STAGE_BYTES + 128);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * C::STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
  // split-K (decode GEMMs, one M tile and few N tiles): gridDim.z CTAs share an output tile, each takes a
  // contiguous range of k-blocks and adds its partial sum into the fp32 output with vector reductions
  const int KB = K / BK / (int)gridDim.z;
  const int kb0 = (int)blockIdx.z * KB;

  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
    if (SPLIT) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAlo) : "memory");
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  pdl_launch_dependents();   // no-ops unless launched with the programmatic-dependent-launch attribute (wide decode)
  if (warp == 0) {
    if (lane == 0) {
      // weights never depend on the previous kernel: the first ring of W tiles is requested before the
      // programmatic-dependency wait, the activation tiles after it
      const int pre = KB < C::STAGES ? KB : C::STAGES;
      for (int kb = 0; kb < pre; ++kb) {
        mbar_expect_tx(full_bar(kb), C::STAGE_BYTES);
        tma_load_2d(base + kb * C::STAGE_BYTES + C::A_TILES * C::A_BYTES, &tmW, (kb0 + kb) * BK, n0, full_bar(kb));
      }
      pdl_wait();
      for (int kb = 0; kb < pre; ++kb) {
        const uint32_t sa = base + kb * C::STAGE_BYTES;
        tma_load_2d(sa, &tmA, (kb0 + kb) * BK, m0, full_bar(kb));
        if (SPLIT) tma_load_2d(sa + C::A_BYTES, &tmAlo, (kb0 + kb) * BK, m0, full_bar(kb));
      }
      for (int kb = pre; kb < KB; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_expect_tx(full_bar(s), C::STAGE_BYTES);
        const uint32_t sa = base + s * C::STAGE_BYTES;
        tma_load_2d(sa, &tmA, (kb0 + kb) * BK, m0, full_bar(s));
        if (SPLIT) tma_load_2d(sa + C::A_BYTES, &tmAlo, (kb0 + kb) * BK, m0, full_bar(s));
        tma_load_2d(sa + C::A_TILES * C::A_BYTES, &tmW, (kb0 + kb) * BK, n0, full_bar(s));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BM, BN);
      for (int kb = 0; kb < KB; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t sa = base + s * C::STAGE_BYTES;
        const uint32_t sb = sa + C::A_TILES * C::A_BYTES;
#pragma unroll
        for (int h = 0; h < C::A_TILES; ++h) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = make_smem_desc(sa + h * C::A_BYTES + k * UMMA_K * 2);
            const uint64_t db = make_smem_desc(sb + k * UMMA_K * 2);
            tc_mma_bf16(tmem_base, da, db, idesc, (kb | k | h) != 0 ? 1u : 0u);
          }
        }
        tc_commit(empty_bar(s));  // slot reusable once these MMAs have read it
      }
      tc_commit(tmem_full_bar);   // accumulator complete
    }
  } else {
    // ---- epilogue: warp w may only touch TMEM lanes [32*(w%4), +32)
    const int quad = warp & 3;
    const int row = m0 + quad * 32 + lane;
    pdl_wait();
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const uint32_t tq = tmem_base + ((uint32_t)(quad * 32) << 16);
    if (EPI == GEMM_SWIGLU_BF16) {
      uint16_t* o = reinterpret_cast<uint16_t*>(out) + (size_t)row * (N / 2) + n0 / 2;
      uint16_t* ol = out_lo ? reinterpret_cast<uint16_t*>(out_lo) + (size_t)row * (N / 2) + n0 / 2 : nullptr;
#pragma unroll 1
      for (int blk = 0; blk < BN / 128; ++blk)
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          uint32_t g[32], u[32];
          tc_ld32(tq + blk * 128 + half * 32, g);
          tc_ld32(tq + blk * 128 + 64 + half * 32, u);
          tc_wait_ld();
          if (row < T) {
            uint32_t pk[16], pl[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float g0 = __uint_as_float(g[2 * i]), g1 = __uint_as_float(g[2 * i + 1]);
              const float a0 = g0 / (1.0f + __expf(-g0)) * __uint_as_float(u[2 * i]);
              const float a1 = g1 / (1.0f + __expf(-g1)) * __uint_as_float(u[2 * i + 1]);
              pk[i] = pack_bf16x2(a0, a1);
              pl[i] = pack_bf16x2(a0 - bf16lo(pk[i]), a1 - bf16hi(pk[i]));
            }
            uint4* dst = reinterpret_cast<uint4*>(o + blk * 64 + half * 32);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            if (ol) {
              uint4* dl = reinterpret_cast<uint4*>(ol + blk * 64 + half * 32);
#pragma unroll
              for (int i = 0; i < 4; ++i) dl[i] = make_uint4(pl[4 * i], pl[4 * i + 1], pl[4 * i + 2], pl[4 * i + 3]);
            }
          }
        }
    } else {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tc_ld32(tq + c * 32, v);
        tc_wait_ld();
        if (row < T) {
          const size_t off = (size_t)row * N + n0 + c * 32;
          if (EPI == GEMM_STORE_BF16) {
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(out) + off);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint32_t pk[4];
#pragma unroll
              for (int j = 0; j < 4; ++j)
                pk[j] = pack_bf16x2(__uint_as_float(v[8 * i + 2 * j]), __uint_as_float(v[8 * i + 2 * j + 1]));
              dst[i] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
          } else {
            float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + off);
            if (gridDim.z > 1) {   // partial sum: out (zeroed by the launcher, or the residual) += acc
#pragma unroll
              for (int i = 0; i < 8; ++i)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + i),
                             "f"(__uint_as_float(v[4 * i])), "f"(__uint_as_float(v[4 * i + 1])),
                             "f"(__uint_as_float(v[4 * i + 2])), "f"(__uint_as_float(v[4 * i + 3]))
                             : "memory");
              continue;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float4 r = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                     __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
              if (EPI == GEMM_RESID_F32) {
                const float4 old = dst[i];
                r.x += old.x; r.y += old.y; r.z += old.z; r.w += old.w;
              }
              dst[i] = r;
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
  }
}

template <int BN, int EPI, bool SPLIT>
cudaError_t launch_tile(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int T,
                        int N, int K, cudaStream_t s, int ksplit = 1, bool pdl = false) {
  CUtensorMap ma, ml, mw;
  if (!make_map(&ma, A, (uint64_t)T, (uint64_t)K, BM) || !make_map(&mw, W, (uint64_t)N, (uint64_t)K, BN) ||
      !make_map(&ml, SPLIT ? Alo : A, (uint64_t)T, (uint64_t)K, BM))
    return cudaErrorInvalidValue;
  auto kern = gemm_tc_kernel<BN, EPI, SPLIT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TileCfg<BN, SPLIT>::SMEM);
  if (e != cudaSuccess) return e;
  dim3 grid(N / BN, (T + BM - 1) / BM, ksplit);
  if (ksplit > 1 && EPI == GEMM_STORE_F32) {
    e = cudaMemsetAsync(out, 0, (size_t)T * N * sizeof(float), s);
    if (e != cudaSuccess) return e;
  }
  // decode-shaped calls (ksplit allowed) are part of a programmatic-dependent-launch chain unless a memset precedes
  return launch_maybe_pdl(kern, grid, dim3(GEMM_THREADS), (size_t)TileCfg<BN, SPLIT>::SMEM, s,
                          pdl && !(ksplit > 1 && EPI == GEMM_STORE_F32), ma, ml, mw, out, out_lo, T, N, K);
}

// largest power-of-two split of the k-blocks that keeps the grid within about one wave and >= 4 k-blocks per CTA
int pick_ksplit(int tiles, int kblocks) {
  int s = 1;
  while (s < 32 && tiles * (s * 2) <= 160 && kblocks % (s * 2) == 0 && kblocks / (s * 2) >= 4) s *= 2;
  return s;
}

template <int EPI, bool SPLIT>
cudaError_t launch_epi(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int T,
                       int N, int K, cudaStream_t s, bool allow_ksplit, bool pdl) {
  // pick the widest tile that still yields at least ~one wave of CTAs on 148 SMs
  const int mt = (T + BM - 1) / BM;
  if constexpr (EPI == GEMM_STORE_F32 || EPI == GEMM_RESID_F32) {
    // decode-shaped call (one M tile): wide tiles + split-K instead of narrow tiles, so that the activations
    // are re-read from L2 by few CTAs and every SM still has work
    if (allow_ksplit && mt == 1 && N % 128 == 0 && N / 128 < 148) {
      const int ks = pick_ksplit(N / 128, K / BK);
      if (ks > 1) return launch_tile<128, EPI, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, ks, pdl);
    }
  }
  if (N % 256 == 0 && (N / 256) * mt >= 148) return launch_tile<256, EPI, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, 1, pdl);
  if constexpr (EPI == GEMM_SWIGLU_BF16) {
    if (N % 128 != 0) return cudaErrorInvalidValue;
    return launch_tile<128, EPI, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, 1, pdl);
  } else {
    if (N % 128 == 0 && (N / 128) * mt >= 148) return launch_tile<128, EPI, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, 1, pdl);
    return launch_tile<64, EPI, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, 1, pdl);
  }
}

template <bool SPLIT>
cudaError_t launch_split(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int T,
                         int N, int K, int epi, cudaStream_t s, bool ks, bool pdl) {
  switch (epi) {
    case GEMM_STORE_BF16: return launch_epi<GEMM_STORE_BF16, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, ks, pdl);
    case GEMM_STORE_F32: return launch_epi<GEMM_STORE_F32, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, ks, pdl);
    case GEMM_RESID_F32: return launch_epi<GEMM_RESID_F32, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, ks, pdl);
    case GEMM_SWIGLU_BF16: return launch_epi<GEMM_SWIGLU_BF16, SPLIT>(A, Alo, W, out, out_lo, T, N, K, s, ks, pdl);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

cudaError_t launch_gemm_tcgen05(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo,
                                int T, int N, int K, int epi, cudaStream_t s, bool allow_ksplit, bool pdl) {
  if (K % BK != 0 || T < 1 || N % 64 != 0) return cudaErrorInvalidValue;
  return Alo ? launch_split<true>(A, Alo, W, out, out_lo, T, N, K, epi, s, allow_ksplit, pdl)
             : launch_split<false>(A, Alo, W, out, out_lo, T, N, K, epi, s, allow_ksplit, pdl);
}

}  // namespace llmc
