// init.cu -- deterministic random-init of the model weights, generated in place in HBM.
//
// Weight definition (the specification the CPU oracle restates independently):
//   key   = mix64(seed ^ (GOLDEN * (tensor_id + 1)))
//   z     = mix64(key + GOLDEN * (index + 1)),  index = row * cols + col of the LOGICAL tensor
//   u     = (int32(z >> 40) - 2^23) * 2^-23
//   W     = bf16_rne(u * amp),  amp = float(init_std * sqrt(3))
//   gain  = bf16_rne(1 + 0.1f * u)    (two separately rounded fp32 operations)
// mix64 is the splitmix64 finaliser.  The packed HBM layouts (fused QKV, block-interleaved
// gate/up) only change WHERE a logical element is stored, never its value.
//
// No counterpart in the reference: johnayoung/llm-consensus holds no model weights (its
// providers are remote, /root/reference/internal/provider/openai.go:141-210).
#include "common.cuh"
#include "kernels.h"

namespace llmc {

__host__ __device__ inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static const uint64_t GOLDEN = 0x9E3779B97F4A7C15ull;

__device__ __forceinline__ float rng_unit(uint64_t key, uint64_t idx) {
  uint64_t z = mix64(key + 0x9E3779B97F4A7C15ull * (idx + 1));
  int v = (int)(z >> 40) - (1 << 23);
  return (float)v * (1.0f / 8388608.0f);
}

// mode 0: dst row = row_off + r; mode 1: gate interleave; mode 2: up interleave
__global__ void __launch_bounds__(256) init_matrix_kernel(uint16_t* __restrict__ dst, int64_t rows, int cols,
                                                          int64_t ld, uint64_t key, float amp, int mode,
                                                          int64_t row_off) {
  const int64_t vec_per_row = cols / 8;
  const int64_t total = rows * vec_per_row;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < total; v += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = v / vec_per_row;
    int c = (int)(v % vec_per_row) * 8;
    int64_t dr = mode == 0 ? row_off + r : (mode == 1 ? gu_row_gate((int)r) : gu_row_up((int)r));
    uint64_t idx = (uint64_t)r * (uint64_t)cols + (uint64_t)c;
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float a = __fmul_rn(rng_unit(key, idx + 2 * i), amp);
      float b = __fmul_rn(rng_unit(key, idx + 2 * i + 1), amp);
      o[i] = pack_bf16x2(a, b);
    }
    *reinterpret_cast<uint4*>(dst + dr * ld + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

__global__ void init_gain_kernel(uint16_t* __restrict__ dst, int n, uint64_t key) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    float p = __fmul_rn(0.1f, rng_unit(key, (uint64_t)i));
    dst[i] = f32_to_bf16_bits(__fadd_rn(1.0f, p));
  }
}

static uint64_t rng_key(uint64_t seed, uint32_t tid) { return mix64(seed ^ (GOLDEN * (uint64_t)(tid + 1))); }

cudaError_t launch_init_matrix(uint16_t* dst, int64_t rows, int cols, int64_t ld, uint64_t seed, uint32_t tid,
                               float amp, int mode, int64_t row_off, cudaStream_t s) {
  int64_t total = rows * (cols / 8);
  int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  if (grid < 1) grid = 1;
  init_matrix_kernel<<<grid, 256, 0, s>>>(dst, rows, cols, ld, rng_key(seed, tid), amp, mode, row_off);
  return cudaGetLastError();
}

cudaError_t launch_init_gain(uint16_t* dst, int n, uint64_t seed, uint32_t tid, cudaStream_t s) {
  init_gain_kernel<<<(n + 255) / 256, 256, 0, s>>>(dst, n, rng_key(seed, tid));
  return cudaGetLastError();
}

}  // namespace llmc
