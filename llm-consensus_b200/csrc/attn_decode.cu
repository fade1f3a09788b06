// attn_decode.cu -- single-query (decode) GQA attention over the paged bf16 KV cache.
//
//   out[h] = softmax(q[h] . K[0..ctx) / sqrt(d)) V[0..ctx)        fp32 softmax, online
//
// HBM-bound on the KV read (ctx * kv_tok bytes per sequence per step).
// Grid (nsplit, kv_heads, B): the context of one (sequence, kv head) is split over `nsplit`
// blocks so that batch 1 still fills the 148 SMs; each block stages 32-position K/V tiles in
// shared memory once and its warps serve the q heads of the GQA group from them.  Split
// partials (o, m, l) go to a scratch buffer; the last block to finish a (sequence, kv head)
// -- found with one atomic counter, no atomics on data -- merges the splits.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the
// boundary; the providers behind it are remote HTTP calls).
#include "common.cuh"
#include "kernels.h"
#include "attn_common.cuh"

namespace llmc {

namespace {

constexpr int TP = ATT_TP;   // positions per tile
constexpr int AWARPS = 4;    // warps per block

template <int D>
__global__ void __launch_bounds__(AWARPS * 32) attn_decode_kernel(const AttnDecodeArgs a) {
  constexpr int KSTRIDE = D + 8;  // padded row (bf16 elements): conflict-free 128-bit row reads
  __shared__ __align__(16) uint16_t ks[TP * KSTRIDE];
  __shared__ __align__(16) uint16_t vs[TP * KSTRIDE];
  __shared__ __align__(16) float qs[8 * D];  // up to 8 q heads per kv head
  __shared__ bool s_last;

  pdl_launch_dependents();
  pdl_wait();

  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = a.q_heads / a.kv_heads;
  const int slot = a.slots[b];
  const int ctx = a.seq_len[slot] + a.ctx_extra;
  const int ntiles = (ctx + TP - 1) / TP;
  const int per = (ntiles + a.nsplit - 1) / a.nsplit;
  const int t0 = split * per, t1 = min(ntiles, t0 + per);
  const float scale = rsqrtf((float)D);
  const int QD = a.q_heads * D;

  for (int i = tid; i < G * D; i += AWARPS * 32) qs[i] = a.q[(size_t)b * QD + (size_t)kvh * G * D + i];

  float m[2], l[2], o[2][D / 32];  // a warp serves heads warp, warp+4 (G <= 8)
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    m[j] = -INFINITY;
    l[j] = 0.f;
#pragma unroll
    for (int i = 0; i < D / 32; ++i) o[j][i] = 0.f;
  }

  for (int t = t0; t < t1; ++t) {
    const int p0 = t * TP;
    const int valid = min(TP, ctx - p0);
    const int page = a.block_table[(size_t)slot * a.bt_stride + p0 / a.page_size];
    const size_t base = (((size_t)page * a.kv_heads + kvh) * a.page_size + (p0 % a.page_size)) * D;
    __syncthreads();  // previous tile fully consumed (also orders the qs fill on the first pass)
    for (int c = tid; c < TP * (D / 8); c += AWARPS * 32) {
      const int r = c / (D / 8), cc = c % (D / 8);
      uint4 kk = make_uint4(0, 0, 0, 0), vv = kk;
      if (r < valid) {
        kk = *reinterpret_cast<const uint4*>(a.kpool + base + (size_t)r * D + cc * 8);
        vv = *reinterpret_cast<const uint4*>(a.vpool + base + (size_t)r * D + cc * 8);
      }
      *reinterpret_cast<uint4*>(ks + r * KSTRIDE + cc * 8) = kk;
      *reinterpret_cast<uint4*>(vs + r * KSTRIDE + cc * 8) = vv;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int hg = warp + j * AWARPS;
      if (hg < G) attend_tile<D>(qs + hg * D, ks, vs, KSTRIDE, valid, scale, m[j], l[j], o[j]);
    }
  }

  // ---- write split partials: [b][head][split][D + 2]
  const int PS = D + 2;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int hg = warp + j * AWARPS;
    if (hg < G) {
      const int head = kvh * G + hg;
      float* pp = a.scratch + (((size_t)b * a.q_heads + head) * a.nsplit + split) * PS;
#pragma unroll
      for (int i = 0; i < D / 32; ++i) pp[lane * (D / 32) + i] = o[j][i];
      if (lane == 0) {
        pp[D] = m[j];
        pp[D + 1] = l[j];
      }
    }
  }
  __threadfence();
  __syncthreads();
  unsigned* ctr = a.counters + (size_t)b * a.kv_heads + kvh;
  if (tid == 0) s_last = (atomicAdd(ctr, 1u) == (unsigned)(a.nsplit - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // ---- merge the splits of this (sequence, kv head): thread = (head in group, dim)
  for (int e = tid; e < G * D; e += AWARPS * 32) {
    const int hg = e / D, i = e - hg * D;
    const int head = kvh * G + hg;
    const float* pp = a.scratch + ((size_t)b * a.q_heads + head) * a.nsplit * PS;
    float M = -INFINITY;
    for (int s = 0; s < a.nsplit; ++s) M = fmaxf(M, __ldcg(pp + (size_t)s * PS + D));
    float L = 0.f, O = 0.f;
    for (int s = 0; s < a.nsplit; ++s) {
      const float ms = __ldcg(pp + (size_t)s * PS + D);
      const float w = (ms == -INFINITY) ? 0.f : __expf(ms - M);
      L = fmaf(__ldcg(pp + (size_t)s * PS + D + 1), w, L);
      O = fmaf(__ldcg(pp + (size_t)s * PS + i), w, O);
    }
    a.out[(size_t)b * QD + (size_t)head * D + i] = O / L;
  }
  if (tid == 0) *ctr = 0u;
}

}  // namespace

size_t attn_decode_scratch_floats(int B, int q_heads, int nsplit, int head_dim) {
  return (size_t)B * q_heads * nsplit * (head_dim + 2);
}

int attn_decode_default_nsplit(int B, int kv_heads) {
  int n = (256 + B * kv_heads - 1) / (B * kv_heads);
  return n < 1 ? 1 : (n > 32 ? 32 : n);
}

cudaError_t launch_attn_decode(const AttnDecodeArgs& a, bool pdl, cudaStream_t s) {
  const int G = a.q_heads / a.kv_heads;
  if (a.q_heads % a.kv_heads != 0 || G > 8 || a.page_size % TP != 0 || a.nsplit < 1) return cudaErrorInvalidValue;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a.nsplit, a.kv_heads, a.B);
  cfg.blockDim = dim3(AWARPS * 32);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  switch (a.head_dim) {
    case 32: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<32>, a);
    case 64: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<64>, a);
    case 128: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<128>, a);
  }
  return cudaErrorInvalidValue;
}

}  // namespace llmc
