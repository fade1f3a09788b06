// gemm_simple.cu -- plain shared-memory tiled CUDA-core GEMM, C = A W^T (bf16 in, fp32 acc).
//
// This is NOT the product prefill path (that is gemm_tcgen05.cu).  It exists as the slow,
// obviously-correct GPU comparator for the tensor-core kernel in the kernel-level tests and as
// an explicit bring-up switch (llmc_config.prefill_impl = 1); it is never selected silently.
#include "common.cuh"
#include "kernels.h"

namespace llmc {

namespace {

constexpr int TM = 64, TN = 64, TK = 32;

template <int EPI>
__global__ void __launch_bounds__(256) gemm_simple_kernel(const uint16_t* __restrict__ A, const uint16_t* __restrict__ Alo,
                                                          const uint16_t* __restrict__ W, void* __restrict__ out,
                                                          void* __restrict__ out_lo, int T, int N, int K) {
  constexpr bool SW = (EPI == GEMM_SWIGLU_BF16);
  __shared__ float As[TK][TM + 4];
  __shared__ float Ws[SW ? 2 : 1][TK][TN + 4];
  const int tid = threadIdx.x;
  const int t0 = blockIdx.y * TM, n0 = blockIdx.x * TN;  // for SWIGLU n0 indexes j (output column)
  const int tx = tid % 16, ty = tid / 16;
  float acc[SW ? 2 : 1][4][4] = {};
  for (int k0 = 0; k0 < K; k0 += TK) {
    // 64 rows x 32 k = 2048 elements per operand, 8 per thread
    {
      const int r = tid / 4, kk = (tid % 4) * 8;
      uint4 a = make_uint4(0, 0, 0, 0), al = a;
      if (t0 + r < T) {
        a = *reinterpret_cast<const uint4*>(A + (size_t)(t0 + r) * K + k0 + kk);
        if (Alo) al = *reinterpret_cast<const uint4*>(Alo + (size_t)(t0 + r) * K + k0 + kk);
      }
      const uint32_t av[4] = {a.x, a.y, a.z, a.w}, lv[4] = {al.x, al.y, al.z, al.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        As[kk + 2 * i][r] = bf16lo(av[i]) + bf16lo(lv[i]);
        As[kk + 2 * i + 1][r] = bf16hi(av[i]) + bf16hi(lv[i]);
      }
#pragma unroll
      for (int h = 0; h < (SW ? 2 : 1); ++h) {
        int row;
        if (SW) row = (h == 0) ? gu_row_gate(n0 + r) : gu_row_up(n0 + r);
        else row = n0 + r;
        uint4 w = make_uint4(0, 0, 0, 0);
        if (row < N) w = *reinterpret_cast<const uint4*>(W + (size_t)row * K + k0 + kk);
        const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          Ws[h][kk + 2 * i][r] = bf16lo(wv[i]);
          Ws[h][kk + 2 * i + 1][r] = bf16hi(wv[i]);
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float a[4], w[SW ? 2 : 1][4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int h = 0; h < (SW ? 2 : 1); ++h)
#pragma unroll
        for (int j = 0; j < 4; ++j) w[h][j] = Ws[h][k][tx * 4 + j];
#pragma unroll
      for (int h = 0; h < (SW ? 2 : 1); ++h)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[h][i][j] = fmaf(a[i], w[h][j], acc[h][i][j]);
    }
    __syncthreads();
  }
  const int NO = SW ? N / 2 : N;  // output columns
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = t0 + ty * 4 + i;
    if (t >= T) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= NO) continue;
      const size_t o = (size_t)t * NO + n;
      if (EPI == GEMM_STORE_BF16) reinterpret_cast<uint16_t*>(out)[o] = f32_to_bf16_bits(acc[0][i][j]);
      else if (EPI == GEMM_STORE_F32) reinterpret_cast<float*>(out)[o] = acc[0][i][j];
      else if (EPI == GEMM_RESID_F32) reinterpret_cast<float*>(out)[o] += acc[0][i][j];
      else {
        const float g = acc[0][i][j], u = acc[SW ? 1 : 0][i][j];
        const float v = g / (1.0f + __expf(-g)) * u;
        const uint16_t hi = f32_to_bf16_bits(v);
        reinterpret_cast<uint16_t*>(out)[o] = hi;
        if (out_lo) reinterpret_cast<uint16_t*>(out_lo)[o] = f32_to_bf16_bits(v - __uint_as_float((uint32_t)hi << 16));
      }
    }
  }
}

}  // namespace

cudaError_t launch_gemm_simple(const uint16_t* A, const uint16_t* Alo, const uint16_t* W, void* out, void* out_lo,
                               int T, int N, int K, int epi, cudaStream_t s) {
  if (K % TK != 0 || T < 1) return cudaErrorInvalidValue;
  const int NO = epi == GEMM_SWIGLU_BF16 ? N / 2 : N;
  dim3 grid((NO + TN - 1) / TN, (T + TM - 1) / TM);
  switch (epi) {
    case GEMM_STORE_BF16: gemm_simple_kernel<GEMM_STORE_BF16><<<grid, 256, 0, s>>>(A, Alo, W, out, out_lo, T, N, K); break;
    case GEMM_STORE_F32: gemm_simple_kernel<GEMM_STORE_F32><<<grid, 256, 0, s>>>(A, Alo, W, out, out_lo, T, N, K); break;
    case GEMM_RESID_F32: gemm_simple_kernel<GEMM_RESID_F32><<<grid, 256, 0, s>>>(A, Alo, W, out, out_lo, T, N, K); break;
    case GEMM_SWIGLU_BF16: gemm_simple_kernel<GEMM_SWIGLU_BF16><<<grid, 256, 0, s>>>(A, Alo, W, out, out_lo, T, N, K); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace llmc
