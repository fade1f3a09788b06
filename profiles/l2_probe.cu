// l2_probe.cu -- micro-benchmark behind DESIGN.md's "L2 look-ahead" experiments: does data requested with
// cp.async.bulk.prefetch.L2 (or prefetch.global.L2) stay in L2 so that a later TMA bulk copy of the same bytes is
// served from L2, (a) in isolation, (b) while the same SMs stream a buffer much larger than L2?
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o profiles/l2_probe profiles/l2_probe.cu
// run:   profiles/l2_probe            (prints GB/s per scenario; one B200)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int SLOT = 32768, NSLOT = 6;

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
      "{\n\t.reg .pred p;\n\tWAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\tbra WAIT_LOOP;\n\tWAIT_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// mode 0: bulk L2 prefetch of [buf, buf + nslots*SLOT), slots dealt round-robin to the blocks
// mode 1: prefetch.global.L2 per 128-byte line
__global__ void prefetch_kernel(const char* buf, long nslots, int mode) {
  for (long s = blockIdx.x; s < nslots; s += gridDim.x) {
    const char* p = buf + s * SLOT;
    if (mode == 0) {
      if (threadIdx.x == 0) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(SLOT) : "memory");
    } else {
      for (int i = threadIdx.x; i < SLOT / 128; i += blockDim.x) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + i * 128));
    }
  }
}

// Streams nslots slots through a 6-slot shared-memory ring with TMA bulk copies (1 producer thread, 1 consumer warp
// that reads one word per lane and releases the slot).  ahead > 0: a second thread issues bulk L2 prefetches
// `ahead` slots (of this block) in front of the producer.
__global__ void __launch_bounds__(96, 1) stream_kernel(const char* buf, long nslots, int ahead, unsigned* sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bars[2 * NSLOT];
  __shared__ volatile long s_issued;
  const uint32_t ring = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(bars);
  if (threadIdx.x == 0) {
    for (int i = 0; i < NSLOT; ++i) { mbar_init(bar + 8 * i, 1); mbar_init(bar + 8 * (NSLOT + i), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    s_issued = 0;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long mine = (nslots - blockIdx.x + gridDim.x - 1) / gridDim.x;
  if (warp == 0) {
    if (lane == 0)
      for (long i = 0; i < mine; ++i) {
        const int r = i % NSLOT;
        const long use = i / NSLOT;
        if (use > 0) mbar_wait(bar + 8 * (NSLOT + r), (use - 1) & 1);
        mbar_expect_tx(bar + 8 * r, SLOT);
        bulk_g2s(ring + r * SLOT, buf + (blockIdx.x + i * gridDim.x) * (long)SLOT, SLOT, bar + 8 * r);
        s_issued = i + 1;
      }
  } else if (warp == 1) {
    unsigned acc = 0;
    for (long i = 0; i < mine; ++i) {
      const int r = i % NSLOT;
      mbar_wait(bar + 8 * r, (i / NSLOT) & 1);
      acc += *reinterpret_cast<volatile unsigned*>(smem + r * SLOT + lane * 4);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + 8 * (NSLOT + r));
    }
    if (acc == 0x12345678u) sink[0] = acc;
  } else if (ahead > 0 && lane == 0) {
    for (long i = 0; i < mine; ++i) {
      while (i >= s_issued + ahead + NSLOT) __nanosleep(100);
      if (i < s_issued + NSLOT) continue;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(buf + (blockIdx.x + i * gridDim.x) * (long)SLOT),
                   "r"(SLOT) : "memory");
    }
  }
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t big = (size_t)3 << 30;
  char* buf;
  unsigned* sink;
  CK(cudaMalloc(&buf, big));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(buf, 1, big));
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NSLOT * SLOT));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  auto flush = [&]() {   // evict: stream the last GB of the buffer
    stream_kernel<<<sms, 96, NSLOT * SLOT>>>(buf + ((size_t)2 << 30), ((size_t)1 << 30) / SLOT, 0, sink);
    CK(cudaDeviceSynchronize());
  };
  auto timed = [&](long nslots, int ahead) {
    CK(cudaEventRecord(e0));
    stream_kernel<<<sms, 96, NSLOT * SLOT>>>(buf, nslots, ahead, sink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return (double)nslots * SLOT / ms / 1e6;
  };
  printf("SMs %d; TMA bulk streaming through a 6 x 32 KB ring per SM, GB/s\n", sms);
  for (int mb : {8, 32, 64, 96, 128, 256}) {
    const long nslots = (long)mb * 1024 * 1024 / SLOT;
    flush();
    const double cold = timed(nslots, 0);
    const double warm = timed(nslots, 0);          // second pass right away: whatever L2 kept
    flush();
    prefetch_kernel<<<sms, 128>>>(buf, nslots, 0);
    CK(cudaDeviceSynchronize());
    const double pf_bulk = timed(nslots, 0);
    flush();
    prefetch_kernel<<<sms, 128>>>(buf, nslots, 1);
    CK(cudaDeviceSynchronize());
    const double pf_line = timed(nslots, 0);
    printf("%4d MB: cold %7.0f | re-read %7.0f | after bulk L2 prefetch %7.0f | after per-line L2 prefetch %7.0f\n", mb, cold,
           warm, pf_bulk, pf_line);
  }
  printf("streaming 2 GB (>> L2) with an in-kernel bulk-prefetch thread `ahead` slots in front of the producer:\n");
  for (int ahead : {0, 2, 4, 8, 16, 32}) {
    flush();
    const double bw = timed(((size_t)2 << 30) / SLOT, ahead);
    printf("  ahead %2d slots/SM (%5.1f MB in L2): %7.0f GB/s\n", ahead, ahead * (double)SLOT * sms / 1e6, bw);
  }
  return 0;
}
