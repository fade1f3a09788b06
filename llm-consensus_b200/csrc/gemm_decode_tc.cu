// gemm_decode_tc.cu -- the matrix products of a wide-batch decode step (4..128 sequences) on the tcgen05
// tensor cores with the operands SWAPPED:   C^T[N][B] = W[N][K] * A[B][K]^T.
//
// HBM-bound (every weight byte is read once per step).  The prefill kernel
// (gemm_tcgen05.cu) makes the activations the 128-row M tile; with B <= 128 rows that wastes its shared
// memory on a mostly empty A tile and leaves only 64 KB of weights in flight per SM -- not enough to cover
// the HBM latency-bandwidth product.  Here the WEIGHTS are the M operand (128 rows x 64 k = 16 KB per
// k-step) and the batch is the N operand (padded to 16/32/64/128), so a pipeline stage is 16 KB of weights +
// 2-32 KB of activations (from L2) and 4-9 stages keep 64-144 KB of the weight stream in flight per SM; the
// MMA work scales with the batch instead of with 128.
//   warp 0      TMA producer (weights tile + activation hi/lo tiles per k-step, 128B swizzle, mbarrier ring)
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer (M=128, N=batch pad, K=16; hi and lo passes)
//   warps 2-5   epilogue: a TMEM lane is a weight row n, a column is a sequence b -> out[b][n]
//               (coalesced over the lanes); fp32 store / residual add / split-K reduction / SwiGLU (gate rows
//               live in lanes 0-63 of the tile, up rows in lanes 64-127: exchanged through shared memory)
// split-K (gridDim.z) as in gemm_tcgen05.cu for the matrices with few row tiles (QKV, O, down).
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "tc_common.cuh"

namespace llmc {

namespace {

using namespace tc;

constexpr int DM = 128, DK = 64, D_UMMA_K = 16;
constexpr int DTHREADS = 192;

__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}

// HALF: half the shared-memory budget per block, so that two blocks fit on an SM -- under programmatic dependent launch
// the NEXT kernel's blocks can then become resident and request their first weight tiles while this kernel's blocks are
// still in their epilogue
template <int NB, bool SPLIT, bool HALF>
struct DCfg {
  static constexpr int W_BYTES = DM * DK * 2;                 // 16 KB
  static constexpr int A_BYTES = NB * DK * 2;                 // NB * 128 B
  static constexpr int A_TILES = SPLIT ? 2 : 1;
  static constexpr int A_ALLOC = (A_BYTES + 1023) / 1024 * 1024;   // swizzled tiles start on 1024 B
  static constexpr int STAGE_BYTES = W_BYTES + A_TILES * A_ALLOC;
  static constexpr int TX_BYTES = W_BYTES + A_TILES * A_BYTES;
  static constexpr int BUDGET = HALF ? 100 * 1024 : 192 * 1024;
  static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 10 ? 10 : (BUDGET / STAGE_BYTES);
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;
  static constexpr int TMEM_COLS = NB < 32 ? 32 : NB;
};

template <int NB, int EPI, bool SPLIT, bool HALF>
__global__ void __launch_bounds__(DTHREADS, 1)
    gemm_decode_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmAlo,
                       const __grid_constant__ CUtensorMap tmW, void* __restrict__ out, void* __restrict__ out_lo,
                       int B, int N, int K) {
  using C = DCfg<NB, SPLIT, HALF>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* base_ptr = smem_raw + (base - raw);
  const uint32_t bar_base = base + C::STAGES * C::STAGE_BYTES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(base_ptr + C::STAGES * C::STAGE_BYTES + 192);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * C::STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * DM;
  const int KB = K / DK / (int)gridDim.z;
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
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
  pdl_launch_dependents();   // the next kernel of the step may start its own prologue

  if (warp == 0) {
    if (lane == 0) {
      // Weights never depend on the previous kernel: the first ring of weight tiles is requested BEFORE the
      // programmatic-dependency wait, the activation tiles (written by the previous kernel) after it.
      const int pre = KB < C::STAGES ? KB : C::STAGES;
      for (int kb = 0; kb < pre; ++kb) {
        mbar_expect_tx(full_bar(kb), C::TX_BYTES);
        tma_load_2d(base + kb * C::STAGE_BYTES, &tmW, (kb0 + kb) * DK, n0, full_bar(kb));
      }
      pdl_wait();
      for (int kb = 0; kb < pre; ++kb) {
        const uint32_t sw = base + kb * C::STAGE_BYTES;
        tma_load_2d(sw + C::W_BYTES, &tmA, (kb0 + kb) * DK, 0, full_bar(kb));
        if (SPLIT) tma_load_2d(sw + C::W_BYTES + C::A_ALLOC, &tmAlo, (kb0 + kb) * DK, 0, full_bar(kb));
      }
      for (int kb = pre; kb < KB; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_expect_tx(full_bar(s), C::TX_BYTES);
        const uint32_t sw = base + s * C::STAGE_BYTES;
        tma_load_2d(sw, &tmW, (kb0 + kb) * DK, n0, full_bar(s));
        tma_load_2d(sw + C::W_BYTES, &tmA, (kb0 + kb) * DK, 0, full_bar(s));
        if (SPLIT) tma_load_2d(sw + C::W_BYTES + C::A_ALLOC, &tmAlo, (kb0 + kb) * DK, 0, full_bar(s));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(DM, NB);
      for (int kb = 0; kb < KB; ++kb) {
        const int s = kb % C::STAGES;
        const uint32_t ph = (kb / C::STAGES) & 1;
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t sw = base + s * C::STAGE_BYTES;
        const uint32_t sa = sw + C::W_BYTES;
#pragma unroll
        for (int h = 0; h < C::A_TILES; ++h) {
#pragma unroll
          for (int k = 0; k < DK / D_UMMA_K; ++k) {
            const uint64_t dw = make_smem_desc(sw + k * D_UMMA_K * 2);
            const uint64_t da = make_smem_desc(sa + h * C::A_ALLOC + k * D_UMMA_K * 2);
            tc_mma_bf16(tmem_base, dw, da, idesc, (kb | k | h) != 0 ? 1u : 0u);
          }
        }
        tc_commit(empty_bar(s));
      }
      tc_commit(tmem_full_bar);
    }
  } else {
    // ---- epilogue: warp w may only touch TMEM lanes [32*(w%4), +32) = weight rows n0 + 32*(w%4) + lane
    const int quad = warp & 3;
    const int r = quad * 32 + lane;        // row inside the tile
    pdl_wait();                            // outputs (and the residual read) must follow the previous kernel
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const uint32_t tq = tmem_base + ((uint32_t)(quad * 32) << 16);
    if (EPI == GEMM_SWIGLU_BF16) {
      // rows 0-63 of the tile: gate j0..j0+63, rows 64-127: the matching up rows (interleaved layout of common.cuh).
      // All MMAs have completed, so the pipeline's shared memory is free: the up half is handed over through it.
      float* xch = reinterpret_cast<float*>(base_ptr);   // [64][NB + 1]
      const int I = N / 2;
      const int j = (n0 / 128) * 64 + (r & 63);
#pragma unroll 1
      for (int c = 0; c < NB / 16; ++c) {
        uint32_t v[16];
        tc_ld16(tq + c * 16, v);
        tc_wait_ld();
        if (quad >= 2) {
#pragma unroll
          for (int i = 0; i < 16; ++i) xch[(r & 63) * (NB + 1) + c * 16 + i] = __uint_as_float(v[i]);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");   // the 4 epilogue warps
        if (quad < 2) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int b = c * 16 + i;
            if (b < B) {
              const float g = __uint_as_float(v[i]);
              const float u = xch[r * (NB + 1) + b];
              const float a = g / (1.0f + __expf(-g)) * u;
              const uint16_t h = f32_to_bf16_bits(a);
              reinterpret_cast<uint16_t*>(out)[(size_t)b * I + j] = h;
              if (out_lo)
                reinterpret_cast<uint16_t*>(out_lo)[(size_t)b * I + j] =
                    f32_to_bf16_bits(a - __uint_as_float((uint32_t)h << 16));
            }
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");   // xch is rewritten by the next chunk
      }
    } else {
      float* o = reinterpret_cast<float*>(out) + n0 + r;
#pragma unroll 1
      for (int c = 0; c < NB / 16; ++c) {
        uint32_t v[16];
        tc_ld16(tq + c * 16, v);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int b = c * 16 + i;
          if (b < B) {
            float* dst = o + (size_t)b * N;
            const float a = __uint_as_float(v[i]);
            if (gridDim.z > 1) atomicAdd(dst, a);                    // partial sum (RED): out zeroed / residual
            else if (EPI == GEMM_RESID_F32) *dst = *dst + a;
            else *dst = a;
          }
        }
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

int pick_ksplit(int tiles, int kblocks, int max_blocks) {
  int s = 1;
  while (s < 32 && tiles * (s * 2) <= max_blocks && kblocks % (s * 2) == 0 && kblocks / (s * 2) >= 4) s *= 2;
  return s;
}

template <int NB, int EPI, bool SPLIT, bool HALF>
cudaError_t launch_nbh(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int B, int N,
                       int K, cudaStream_t s, bool pdl) {
  using C = DCfg<NB, SPLIT, HALF>;
  CUtensorMap ma, ml, mw;
  if (!make_map(&ma, A, (uint64_t)B, (uint64_t)K, NB) || !make_map(&mw, W, (uint64_t)N, (uint64_t)K, DM) ||
      !make_map(&ml, SPLIT ? Alo : A, (uint64_t)B, (uint64_t)K, NB))
    return cudaErrorInvalidValue;
  auto kern = gemm_decode_kernel<NB, EPI, SPLIT, HALF>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
  if (e != cudaSuccess) return e;
  const int tiles = N / DM;
  const int ks = (EPI == GEMM_SWIGLU_BF16) ? 1 : pick_ksplit(tiles, K / DK, HALF ? 300 : 160);
  if (ks > 1 && EPI == GEMM_STORE_F32) {
    e = cudaMemsetAsync(out, 0, (size_t)B * N * sizeof(float), s);
    if (e != cudaSuccess) return e;
  }
  // (a memset node in front of the kernel is an ordinary dependency: no programmatic edge then)
  return launch_maybe_pdl(kern, dim3(tiles, 1, ks), dim3(DTHREADS), (size_t)C::SMEM, s,
                          pdl && !(ks > 1 && EPI == GEMM_STORE_F32), ma, ml, mw, out, out_lo, B, N, K);
}

template <int NB, int EPI, bool SPLIT>
cudaError_t launch_nb(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int B, int N,
                      int K, cudaStream_t s, bool pdl) {
  // batch <= 32: two blocks per SM (measured at the 1B shape: batch 8 1.044 -> 0.955 ms per step, 16: 1.086 -> 0.983,
  // 32: 1.386 -> 1.322); LLMC_GEMM_HALF=0 restores one block per SM with the whole shared memory as its ring
  static const bool half = !(getenv("LLMC_GEMM_HALF") && atoi(getenv("LLMC_GEMM_HALF")) == 0);
  if constexpr (NB <= 32) {
    if (half) return launch_nbh<NB, EPI, SPLIT, true>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
  }
  return launch_nbh<NB, EPI, SPLIT, false>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
}

template <int EPI, bool SPLIT>
cudaError_t launch_b(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int B, int N,
                     int K, cudaStream_t s, bool pdl) {
  if (B <= 16) return launch_nb<16, EPI, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
  if (B <= 32) return launch_nb<32, EPI, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
  if (B <= 64) return launch_nb<64, EPI, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
  return launch_nb<128, EPI, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
}

template <bool SPLIT>
cudaError_t launch_e(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo, int B, int N,
                     int K, int epi, cudaStream_t s, bool pdl) {
  switch (epi) {
    case GEMM_STORE_F32: return launch_b<GEMM_STORE_F32, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
    case GEMM_RESID_F32: return launch_b<GEMM_RESID_F32, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
    case GEMM_SWIGLU_BF16: return launch_b<GEMM_SWIGLU_BF16, SPLIT>(A, Alo, W, out, out_lo, B, N, K, s, pdl);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

bool gemm_decode_tc_supported(int B, int N, int K, int epi) {
  return B >= 1 && B <= 128 && N % DM == 0 && K % DK == 0 &&
         (epi == GEMM_STORE_F32 || epi == GEMM_RESID_F32 || epi == GEMM_SWIGLU_BF16);
}

cudaError_t launch_gemm_decode_tc(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo,
                                  int B, int N, int K, int epi, cudaStream_t s, bool pdl) {
  if (!gemm_decode_tc_supported(B, N, K, epi)) return cudaErrorInvalidValue;
  return Alo ? launch_e<true>(A, Alo, W, out, out_lo, B, N, K, epi, s, pdl)
             : launch_e<false>(A, Alo, W, out, out_lo, B, N, K, epi, s, pdl);
}

}  // namespace llmc
