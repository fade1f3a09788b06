// common.cuh -- shared device/host helpers for the llmc CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/llmc.h"

namespace llmc {

// ---------------------------------------------------------------- model description
struct ModelDims {
  int hidden, layers, q_heads, kv_heads, head_dim, inter, vocab, tied;
  float rope_theta, rms_eps, init_std;
  int qd() const { return q_heads * head_dim; }
  int kd() const { return kv_heads * head_dim; }
  int qkv_rows() const { return qd() + 2 * kd(); }
};

// tensor ids of the weight RNG (specification shared with the oracle, see init.cu)
enum { K_EMB = 0, K_LMHEAD = 1, K_FINAL_NORM = 2 };
enum { K_ATTN_NORM = 0, K_WQ, K_WK, K_WV, K_WO, K_MLP_NORM, K_WG, K_WU, K_WD };
static inline uint32_t tensor_id(int layer, int kind) { return (uint32_t)((layer + 1) * 16 + kind); }

// gate/up rows are interleaved in blocks of GU_BLK so that one 128-wide GEMM tile holds
// matching gate and up columns: packed row of gate j = (j/64)*128 + j%64, up j = that + 64.
constexpr int GU_BLK = 64;
__host__ __device__ inline int gu_row_gate(int j) { return (j / GU_BLK) * (2 * GU_BLK) + (j % GU_BLK); }
__host__ __device__ inline int gu_row_up(int j) { return gu_row_gate(j) + GU_BLK; }

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint16_t f32_to_bf16_bits(float f) {
  return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  return (uint32_t)f32_to_bf16_bits(lo) | ((uint32_t)f32_to_bf16_bits(hi) << 16);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// streaming 128-bit load of read-once data (weights): bypass L1 allocation
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// Programmatic dependent launch: let the next kernel in the stream start its prologue
// (weight prefetch) early, and wait for the previous kernel's results before touching them.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

#ifdef __CUDACC__
// Launch with (pdl = true) or without the programmatic-dependent-launch attribute.  A kernel launched with it MUST
// execute pdl_wait() before it reads anything its predecessor wrote (and before it writes anything the predecessor
// reads); everything before that point overlaps the predecessor's tail.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_maybe_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl,
                                    Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// ---------------------------------------------------------------- host error plumbing
struct Status {
  int code = LLMC_OK;
  std::string msg;
  bool ok() const { return code == LLMC_OK; }
};

#define LLMC_CUDA_TRY(st, expr)                                                              \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      (st).code = (_e == cudaErrorMemoryAllocation) ? LLMC_ERR_OOM : LLMC_ERR_CUDA;          \
      (st).msg = std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" __FILE__ ":" +    \
                 std::to_string(__LINE__) + ")";                                             \
      return (st).code;                                                                      \
    }                                                                                        \
  } while (0)

}  // namespace llmc
