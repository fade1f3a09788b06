// noalloc.cu -- an LM-head + greedy-argmax entry point for callers that are not allowed to allocate
// device memory.
//
// Who needs it: code that runs this library inside somebody else's memory budget -- a serving
// process whose device memory is carved up front by its own arena or pool allocator, a plugin
// hosted by a framework that owns every allocation, or a real-time loop where cudaMalloc/cudaFree
// (which synchronise the device and take driver locks) must never happen after start-up.  Such a
// caller can hand us buffers it already owns for the inputs and outputs, but it has nowhere to get
// the kernel's private scratch (the per-block argmax candidates and the "last block" counter) from.
// Here that scratch lives in a fixed-size array that is part of the loaded library image, so no
// allocation call of any kind is made, by us or by a library we call.
//
// The price, stated once: the array is sized for the stated maximum (batch 8, 262144 vocabulary
// rows -> 2 MiB + 512 B) and every GPU on which the library is loaded keeps it until the process
// exits; it is shared by all calls, so two calls whose kernels overlap on the same GPU race on it.
#include <mutex>
#include <string>

#include "../../include/llmc_ops.h"
#include "common.cuh"
#include "kernels.h"

using namespace llmc;

namespace {

constexpr int NA_MAX_BATCH = LLMC_NOALLOC_MAX_BATCH;
constexpr int NA_MAX_VOCAB = LLMC_NOALLOC_MAX_VOCAB;
// candidates per batch row the array has room for; every call checks the block count the kernels would
// actually use (gemv_lmhead_blocks / gemv_mma_lmhead_blocks) against it, so a change of their launch
// geometry is refused loudly instead of overrunning the array
constexpr int NA_MAX_BLOCKS = (NA_MAX_VOCAB / 4 + 1) / 2;
// layout (bytes): [0,256) control words cleared before every launch: block counter, step counter,
// pending-token and length cells the kernel's finalisation writes; [256,512) identity slot table,
// written once at set-up; then the per-block argmax candidates (values, then indices)
constexpr size_t NA_CTRL_BYTES = 256, NA_SLOTS_OFF = 256, NA_PART_OFF = 512;
constexpr size_t NA_PART_ELEMS = (size_t)NA_MAX_BATCH * NA_MAX_BLOCKS;
constexpr size_t NA_BYTES = NA_PART_OFF + NA_PART_ELEMS * (sizeof(float) + sizeof(int));

// In host code this name is only a stand-in for the device array, not a device address; the
// address on each GPU comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char* g_ptr[MAX_DEVICES] = {};
thread_local std::string g_err;

int nfail(int code, const std::string& m) {
  g_err = m;
  return code;
}

}  // namespace

extern "C" {

const char* llmc_noalloc_last_error(void) { return g_err.c_str(); }

int llmc_lmhead_argmax_noalloc(int device, int V, int K, int B, const uint16_t* d_W, const float* d_x,
                               const uint16_t* d_gain, float eps, float* d_logits, int32_t* d_tokens,
                               void* stream) {
  if (!d_W || !d_x || !d_gain || !d_tokens) return nfail(LLMC_ERR_INVALID, "llmc_lmhead_argmax_noalloc: null buffer");
  if (B < 1 || B > NA_MAX_BATCH)
    return nfail(LLMC_ERR_LIMIT, "llmc_lmhead_argmax_noalloc: batch " + std::to_string(B) + " is outside [1, " +
                                     std::to_string(NA_MAX_BATCH) + "], the size the built-in scratch was compiled for");
  if (V < 4 || V > NA_MAX_VOCAB || V % 4)
    return nfail(LLMC_ERR_LIMIT, "llmc_lmhead_argmax_noalloc: vocabulary " + std::to_string(V) +
                                     " must be a multiple of 4 no larger than " + std::to_string(NA_MAX_VOCAB) +
                                     ", the size the built-in scratch was compiled for");
  if (K < 256 || K % 256) return nfail(LLMC_ERR_INVALID, "llmc_lmhead_argmax_noalloc: K must be a positive multiple of 256");
  if (device < 0 || device >= MAX_DEVICES) return nfail(LLMC_ERR_INVALID, "llmc_lmhead_argmax_noalloc: bad device ordinal");
  const int need_blocks = B == 1 ? gemv_lmhead_blocks(V) : gemv_mma_lmhead_blocks(V);
  if (need_blocks > NA_MAX_BLOCKS)
    return nfail(LLMC_ERR_LIMIT, "llmc_lmhead_argmax_noalloc: the kernel would use " + std::to_string(need_blocks) +
                                     " blocks, the built-in scratch holds candidates of " + std::to_string(NA_MAX_BLOCKS));
  // the caller's current device is put back before every return
  int prev_device = -1;
  cudaError_t ce = cudaGetDevice(&prev_device);
  if (ce != cudaSuccess) return nfail(LLMC_ERR_CUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(ce));
  struct Restore {
    int dev;
    ~Restore() { if (dev >= 0) cudaSetDevice(dev); }
  } restore{prev_device == device ? -1 : prev_device};
  ce = cudaSetDevice(device);
  if (ce != cudaSuccess) return nfail(LLMC_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(ce));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);

  // ---- set-up, once per device, with that device current: resolve the array's address ON THIS GPU
  unsigned char* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ptr[device]) {
      unsigned char* p = nullptr;
      ce = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
      if (ce != cudaSuccess || !p)
        return nfail(LLMC_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(ce));
      int ident[NA_MAX_BATCH];
      for (int i = 0; i < NA_MAX_BATCH; ++i) ident[i] = i;
      // on the caller's stream, ahead of the launch below (a pageable source is staged before the call returns)
      ce = cudaMemcpyAsync(p + NA_SLOTS_OFF, ident, sizeof(ident), cudaMemcpyHostToDevice, s);
      if (ce != cudaSuccess) return nfail(LLMC_ERR_CUDA, std::string("scratch set-up copy: ") + cudaGetErrorString(ce));
      g_ptr[device] = p;
    }
    base = g_ptr[device];
  }

  // the array keeps whatever the previous call left: everything the kernel expects to find at zero
  // (block counter, step counter) is cleared on the caller's stream before the launch
  ce = cudaMemsetAsync(base, 0, NA_CTRL_BYTES, s);
  if (ce != cudaSuccess) return nfail(LLMC_ERR_CUDA, std::string("clearing scratch: ") + cudaGetErrorString(ce));

  GemvArgs a;
  memset(&a, 0, sizeof(a));
  a.W = d_W; a.N = V; a.K = K; a.B = B;
  a.x = d_x; a.x_stride = K; a.gain = d_gain; a.eps = eps;
  a.y = d_logits; a.y_stride = V;
  a.slots = reinterpret_cast<const int*>(base + NA_SLOTS_OFF);
  a.counter = reinterpret_cast<unsigned*>(base);
  a.step_counter = reinterpret_cast<int*>(base + 4);
  a.next_token_out = reinterpret_cast<int*>(base + 64);
  a.seq_len_out = reinterpret_cast<int*>(base + 128);
  a.seq_inc = 0;
  a.part_val = reinterpret_cast<float*>(base + NA_PART_OFF);
  a.part_idx = reinterpret_cast<int*>(base + NA_PART_OFF + NA_PART_ELEMS * sizeof(float));
  a.out_tokens = d_tokens;  // step counter is 0, stride 1: token of batch row b lands in d_tokens[b]
  a.out_stride = 1;
  ce = launch_gemv_any(EPI_LMHEAD, PRO_NORM, a, false, s);
  if (ce != cudaSuccess) return nfail(LLMC_ERR_CUDA, std::string("launch: ") + cudaGetErrorString(ce));
  return LLMC_OK;
}

}  // extern "C"
