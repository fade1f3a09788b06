// ring_common.cuh -- pieces shared by the two shared-memory-ring decode kernels (decode_ring.cu: phases separated
// by grid barriers; decode_flow.cu: phases chained by tagged words): ring geometry, the phase descriptor, mbarrier
// and TMA bulk-copy wrappers.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace llmc {
namespace ring {

constexpr int RW = 8;                  // consumer warps: two groups of 4
constexpr int RTHREADS = (RW + 4) * 32;   // two consumer warpgroups + one producer warpgroup (one warp of it works)
constexpr int SEG = 2048;              // k elements per warp segment
constexpr int SLOT_BYTES = 32768;      // 8 row segments of SEG bf16
constexpr int NSLOT = 6;
constexpr int RING_MAX_SPLIT = 16;     // blocks per kv head in the attention phase (scratch layout [kv head][16][G])
enum { MODE_QKV = 0, MODE_RESID = 1, MODE_SWIGLU = 2, MODE_LMHEAD = 3 };

struct RPhase {
  const uint16_t* W;
  int ntasks;  // row pairs
  int K;
  int mode;
  int hd2;     // MODE_QKV: head_dim / 2
};

__device__ __forceinline__ void pair_rows(const RPhase& p, int g, int& r0, int& r1) {
  if (p.mode == MODE_QKV) {
    const int h = g / p.hd2, i = g - h * p.hd2;
    r0 = h * 2 * p.hd2 + i;
    r1 = r0 + p.hd2;
  } else if (p.mode == MODE_SWIGLU) {
    r0 = gu_row_gate(g);
    r1 = gu_row_up(g);
  } else {
    r0 = 2 * g;
    r1 = 2 * g + 1;
  }
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
// barriers of the consumer warps only (the producer warp never joins a block-wide barrier after start-up)
__device__ __forceinline__ void cbar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ void gbar(int group) { asm volatile("bar.sync %0, 128;" ::"r"(2 + group) : "memory"); }

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

}  // namespace ring
}  // namespace llmc
