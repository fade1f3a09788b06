// gemv.cu -- the decode-step matrix-vector family: y = W x with bf16 weights streamed once
// from HBM, fp32 activations and accumulation, and the neighbouring small ops fused in:
//   prologue: RMSNorm (gain folded into x, 1/rms applied to the dot products), embedding gather
//   epilogue: residual add | RoPE + paged-KV append | SwiGLU | greedy argmax (+ step finalise)
//
// HBM-bound (W bytes are read once per decode step).  Design:
//   * one warp owns R rows; each lane streams 16-byte chunks (8 bf16) of those rows with
//     L1::no_allocate loads, two register batches of 8 loads in flight per lane; small matrices
//     additionally split K over 2 or 4 warps of a block so that more warps stream at once;
//   * the first batch of weight loads is issued BEFORE griddepcontrol.wait, so under
//     programmatic dependent launch the weight stream of kernel n+1 overlaps the tail of
//     kernel n (weights never depend on a previous kernel);
//   * x is staged once per block in shared memory, permuted so that lane reads are
//     conflict-free 128-bit loads; warp-shuffle reductions; no atomics on data.
//
// No counterpart in the reference (remote HTTP providers only,
// /root/reference/internal/provider/provider.go:13-21 is the boundary this sits behind).
#include "common.cuh"
#include "kernels.h"

namespace llmc {

namespace {

// A block has WARPS warps = RG row groups x KS k-slices: warp w works on row group w / KS and on
// the k-range slice w % KS of its rows; the KS partial sums meet in shared memory.  Splitting K
// multiplies the number of warps streaming a small matrix (O, QKV, down at batch 1), which is
// what keeps enough bytes in flight to cover the HBM latency.
template <int EPI>
struct GemvCfg {
  static constexpr int R = (EPI == EPI_LMHEAD) ? 4 : 2;       // rows per warp
  static constexpr int WARPS = 8;
};

// smem position of element k of a staged activation row (see file header)
__device__ __forceinline__ int xs_index(int k) {
  int c = k >> 8, w = k & 255;
  int lane = w >> 3, j = w & 7;
  return (c << 8) + ((j >> 2) << 7) + (lane << 2) + (j & 3);
}

__device__ __forceinline__ float dot8(const uint4& w, const float4& a, const float4& b) {
  float s = bf16lo(w.x) * a.x;
  s = fmaf(bf16hi(w.x), a.y, s);
  s = fmaf(bf16lo(w.y), a.z, s);
  s = fmaf(bf16hi(w.y), a.w, s);
  s = fmaf(bf16lo(w.z), b.x, s);
  s = fmaf(bf16hi(w.z), b.y, s);
  s = fmaf(bf16lo(w.w), b.z, s);
  s = fmaf(bf16hi(w.w), b.w, s);
  return s;
}

template <int EPI, int PRO, int BT, int KS>
__global__ void __launch_bounds__(GemvCfg<EPI>::WARPS * 32) gemv_kernel(const GemvArgs a) {
  constexpr int R = GemvCfg<EPI>::R;
  constexpr int WARPS = GemvCfg<EPI>::WARPS;
  constexpr int RG = WARPS / KS;  // row groups per block
  constexpr int UK = 8 / R;       // k-iterations per register batch (8 loads per lane per batch, 2 batches in flight)
  extern __shared__ __align__(16) float smem[];
  float* xs = smem;                        // [BT][K] permuted
  float* red = smem + (size_t)BT * a.K;    // [WARPS][BT] sum of squares
  float* ksum = red + WARPS * BT;          // [WARPS][R][BT] k-slice partial sums
  __shared__ float s_val[WARPS][BT];
  __shared__ int s_idx[WARPS][BT];
  __shared__ bool s_last;

  pdl_launch_dependents();

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int K = a.K, NIT = K >> 8;
  const int ngroups = a.N / R;
  const int rg = warp / KS, ks = warp % KS;
  const int g = blockIdx.x * RG + rg;
  const bool active = g < ngroups;
  // this warp's k-iterations: [it_lo, it_hi)
  const int it_per = (NIT + KS - 1) / KS;
  const int it_lo = ks * it_per, it_hi = min(NIT, it_lo + it_per);

  // ---- rows of this warp
  int rows[R];
  if (EPI == EPI_QKV_ROPE) {
    int hd2 = a.head_dim >> 1;
    int h = g / hd2, i = g - h * hd2;
    rows[0] = h * a.head_dim + i;
    rows[1] = rows[0] + hd2;
  } else if (EPI == EPI_SWIGLU) {
    rows[0] = gu_row_gate(g);
    rows[1] = gu_row_up(g);
  } else {
#pragma unroll
    for (int r = 0; r < R; ++r) rows[r] = g * R + r;
  }
  const uint16_t* wp[R];
#pragma unroll
  for (int r = 0; r < R; ++r) wp[r] = a.W + (size_t)(active ? rows[r] : 0) * K + lane * 8;

  // two register batches of R * UK = 8 loads (128 B per lane, 4 KB per warp) each: batch i+1 is
  // requested before batch i is consumed.  (A finer-grained rolling refill measured 2x slower:
  // the 6 scoreboard slots per warp make a consumer wait for the youngest load sharing its slot.)
  uint4 wcur[R][UK], wnxt[R][UK];
  auto load_batch = [&](uint4(&w)[R][UK], int it0) {
#pragma unroll
    for (int u = 0; u < UK; ++u) {
      const int it = it0 + u;
#pragma unroll
      for (int r = 0; r < R; ++r)
        w[r][u] = (active && it < it_hi) ? ldg_stream(wp[r] + (size_t)it * 256) : make_uint4(0, 0, 0, 0);
    }
  };
  // weights do not depend on the previous kernel: start streaming before the dependency wait
  load_batch(wcur, it_lo);
  load_batch(wnxt, it_lo + UK);

  pdl_wait();

  // epilogue metadata of batch row `lane`, fetched now so that its dependent-load chain
  // (slot -> position -> page / rope entry) overlaps the weight stream instead of trailing it
  int e_pos = 0, e_page = 0;
  float e_cos = 1.f, e_sin = 0.f;
  if (EPI == EPI_QKV_ROPE && active && ks == 0 && lane < a.B) {
    const int hd2 = a.head_dim >> 1;
    const int slot = a.slots[lane];
    e_pos = a.seq_len[slot];
    e_page = a.block_table[(size_t)slot * a.bt_stride + e_pos / a.page_size];
    e_cos = a.rope_cos[(size_t)e_pos * hd2 + (g % hd2)];
    e_sin = a.rope_sin[(size_t)e_pos * hd2 + (g % hd2)];
  }

  // ---- prologue: stage x (and fold the RMSNorm gain), sum of squares per batch row
  float ss[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) ss[b] = 0.f;
  for (int k = tid * 4; k < K; k += WARPS * 32 * 4) {
    float4 gv = make_float4(1.f, 1.f, 1.f, 1.f);
    if (PRO != PRO_PLAIN) {
      uint2 gb = *reinterpret_cast<const uint2*>(a.gain + k);
      gv = make_float4(bf16lo(gb.x), bf16hi(gb.x), bf16lo(gb.y), bf16hi(gb.y));
    }
    const int si = xs_index(k);
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b < a.B) {
        if (PRO == PRO_EMBED_NORM) {
          int tok = a.next_token[a.slots[b]];
          uint2 e = *reinterpret_cast<const uint2*>(a.emb + (size_t)tok * K + k);
          v = make_float4(bf16lo(e.x), bf16hi(e.x), bf16lo(e.y), bf16hi(e.y));
          if (blockIdx.x == 0) *reinterpret_cast<float4*>(a.x_out + (size_t)b * K + k) = v;
        } else {
          v = *reinterpret_cast<const float4*>(a.x + (size_t)b * a.x_stride + k);
        }
      }
      if (PRO != PRO_PLAIN) ss[b] += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
      v.x *= gv.x; v.y *= gv.y; v.z *= gv.z; v.w *= gv.w;
      *reinterpret_cast<float4*>(xs + (size_t)b * K + si) = v;
    }
  }
  if (PRO != PRO_PLAIN) {
#pragma unroll
    for (int b = 0; b < BT; ++b) {
      float t = warp_sum(ss[b]);
      if (lane == 0) red[warp * BT + b] = t;
    }
  }
  __syncthreads();
  float rs[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) {
    rs[b] = 1.f;
    if (PRO != PRO_PLAIN) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < WARPS; ++w) t += red[w * BT + b];
      rs[b] = rsqrtf(t / (float)K + a.eps);
    }
  }

  // ---- main loop: consume batch i while batch i+1 is in flight
  float acc[R][BT];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int b = 0; b < BT; ++b) acc[r][b] = 0.f;

  // ping-pong: the batch just consumed is re-requested at once and left alone until the other batch
  // has been consumed (copying wnxt into wcur would stall on the loads still in flight)
  auto consume = [&](uint4(&w)[R][UK], int it0) {
#pragma unroll
    for (int u = 0; u < UK; ++u) {
      const int it = it0 + u;
      if (it < it_hi) {
        const float* xp = xs + (it << 8) + (lane << 2);
#pragma unroll
        for (int b = 0; b < BT; ++b) {
          const float4 x0 = *reinterpret_cast<const float4*>(xp + (size_t)b * K);
          const float4 x1 = *reinterpret_cast<const float4*>(xp + (size_t)b * K + 128);
#pragma unroll
          for (int r = 0; r < R; ++r) acc[r][b] += dot8(w[r][u], x0, x1);
        }
      }
    }
    if (it0 + 2 * UK < it_hi) load_batch(w, it0 + 2 * UK);
  };
  for (int it0 = it_lo; it0 < it_hi; it0 += 2 * UK) {
    consume(wcur, it0);
    if (it0 + UK < it_hi) consume(wnxt, it0 + UK);
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int b = 0; b < BT; ++b) acc[r][b] = warp_sum(acc[r][b]);
  if (KS > 1) {
    // combine the k-slices: every warp publishes, the slice-0 warp of each row group sums
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int b = 0; b < BT; ++b) ksum[(warp * R + r) * BT + b] = acc[r][b];
    }
    __syncthreads();
    if (ks == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int b = 0; b < BT; ++b) {
          float t = 0.f;
#pragma unroll
          for (int k2 = 0; k2 < KS; ++k2) t += ksum[((warp + k2) * R + r) * BT + b];
          acc[r][b] = t;
        }
    }
  }
  const bool owner = active && ks == 0;   // the warp that finishes this row group
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int b = 0; b < BT; ++b) acc[r][b] *= rs[b];

  // ---- epilogues: after the butterfly every lane holds every sum; lane b finishes batch row b
  if (EPI == EPI_STORE || EPI == EPI_RESID) {
    if (owner && lane < a.B) {
#pragma unroll
      for (int b = 0; b < BT; ++b)
        if (lane == b) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            float* p = a.y + (size_t)b * a.y_stride + rows[r];
            *p = (EPI == EPI_RESID) ? (*p + acc[r][b]) : acc[r][b];
          }
        }
    }
  } else if (EPI == EPI_SWIGLU) {
    if (owner && lane < a.B) {
#pragma unroll
      for (int b = 0; b < BT; ++b)
        if (lane == b) {
          float gt = acc[0][b], up = acc[1][b];
          a.y[(size_t)b * a.y_stride + g] = gt / (1.0f + __expf(-gt)) * up;
        }
    }
  } else if (EPI == EPI_QKV_ROPE) {
    if (owner && lane < a.B) {
      const int d = a.head_dim, hd2 = d >> 1;
      const int h = g / hd2, i = g - h * hd2;
#pragma unroll
      for (int b = 0; b < BT; ++b)
        if (lane == b) {
          const int pos = e_pos;
          float v0 = acc[0][b], v1 = acc[1][b];
          if (h < a.q_heads + a.kv_heads) {
            const float o0 = v0 * e_cos - v1 * e_sin, o1 = v1 * e_cos + v0 * e_sin;
            v0 = o0; v1 = o1;
          }
          if (h < a.q_heads) {
            float* qp = a.y + (size_t)b * a.y_stride + h * d + i;
            qp[0] = v0;
            qp[hd2] = v1;
          } else {
            const bool isk = h < a.q_heads + a.kv_heads;
            const int kvh = isk ? h - a.q_heads : h - a.q_heads - a.kv_heads;
            const int page = e_page;
            const int off = pos % a.page_size;
            const size_t eo = (((size_t)page * a.kv_heads + kvh) * a.page_size + off) * d + i;
            if (a.kv_fp32) {
              float* dst = reinterpret_cast<float*>(isk ? a.kpool : a.vpool) + eo;
              dst[0] = v0;
              dst[hd2] = v1;
            } else {
              uint16_t* dst = reinterpret_cast<uint16_t*>(isk ? a.kpool : a.vpool) + eo;
              dst[0] = f32_to_bf16_bits(v0);
              dst[hd2] = f32_to_bf16_bits(v1);
            }
          }
        }
    }
  } else if (EPI == EPI_LMHEAD) {
    // logits (optional) + greedy argmax; ties resolve to the lowest row index
    if (lane < BT) {
#pragma unroll
      for (int b = 0; b < BT; ++b)
        if (lane == b) {
          float best = -INFINITY;
          int bi = 0x7fffffff;
          if (owner && b < a.B) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
              float v = acc[r][b];
              if (a.y) a.y[(size_t)b * a.y_stride + rows[r]] = v;
              if (v > best) { best = v; bi = rows[r]; }
            }
          }
          s_val[warp][b] = best;
          s_idx[warp][b] = bi;
        }
    }
    __syncthreads();
    const int nblk = gridDim.x;
    if (tid < BT && tid < a.B) {
      float best = s_val[0][tid];
      int bi = s_idx[0][tid];
#pragma unroll
      for (int w = 1; w < WARPS; ++w)
        if (s_val[w][tid] > best) { best = s_val[w][tid]; bi = s_idx[w][tid]; }
      a.part_val[(size_t)tid * nblk + blockIdx.x] = best;
      a.part_idx[(size_t)tid * nblk + blockIdx.x] = bi;
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(a.counter, 1u) == (unsigned)(nblk - 1));
    __syncthreads();
    if (s_last) {
      __threadfence();
      for (int b = 0; b < a.B; ++b) {
        float best = -INFINITY;
        int bi = 0x7fffffff;
        for (int j = tid; j < nblk; j += WARPS * 32) {
          float v = __ldcg(a.part_val + (size_t)b * nblk + j);
          int ix = __ldcg(a.part_idx + (size_t)b * nblk + j);
          if (v > best || (v == best && ix < bi)) { best = v; bi = ix; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          float ov = __shfl_xor_sync(0xffffffffu, best, o);
          int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        __syncthreads();
        if (lane == 0) { s_val[warp][0] = best; s_idx[warp][0] = bi; }
        __syncthreads();
        if (tid == 0) {
          for (int w = 1; w < WARPS; ++w)
            if (s_val[w][0] > best || (s_val[w][0] == best && s_idx[w][0] < bi)) { best = s_val[w][0]; bi = s_idx[w][0]; }
          const int slot = a.slots[b];
          a.next_token_out[slot] = bi;
          a.seq_len_out[slot] += a.seq_inc;
          if (a.out_tokens) a.out_tokens[(size_t)b * a.out_stride + *a.step_counter] = bi;
        }
      }
      if (tid == 0) {
        *a.step_counter += 1;
        *a.counter = 0u;
      }
    }
  }
}

template <int EPI, int PRO, int BT, int KS>
cudaError_t launch_ks(const GemvArgs& a, bool pdl, cudaStream_t s) {
  constexpr int R = GemvCfg<EPI>::R, WARPS = GemvCfg<EPI>::WARPS, RG = WARPS / KS;
  const int ngroups = a.N / R;
  const int grid = (ngroups + RG - 1) / RG;
  const size_t smem = ((size_t)BT * a.K + (size_t)WARPS * BT + (size_t)WARPS * R * BT) * sizeof(float);
  auto kern = gemv_kernel<EPI, PRO, BT, KS>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(WARPS * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, a);
}

// k-slices per row group: enough warps to keep ~4096 row-group tasks in flight on small matrices
int pick_ks(int N, int K, int R) {
  const int tasks = N / R, nit = K >> 8;
  int ks = 1;
  while (ks < 4 && tasks * ks < 4096 && R * nit / (ks * 2) >= 8) ks *= 2;   // keep >= 8 loads per lane
  return ks;
}

template <int EPI, int PRO, int BT>
cudaError_t launch_one(const GemvArgs& a, bool pdl, cudaStream_t s) {
  switch (pick_ks(a.N, a.K, GemvCfg<EPI>::R)) {
    case 1: return launch_ks<EPI, PRO, BT, 1>(a, pdl, s);
    case 2: return launch_ks<EPI, PRO, BT, 2>(a, pdl, s);
    default: return launch_ks<EPI, PRO, BT, 4>(a, pdl, s);
  }
}

template <int EPI, int PRO>
cudaError_t launch_bt(int bt, const GemvArgs& a, bool pdl, cudaStream_t s) {
  // batches of 2..8 rows run on the tensor cores (gemv_mma.cu); this CUDA-core stream serves one row
  if (bt == 1) return launch_one<EPI, PRO, 1>(a, pdl, s);
  return cudaErrorInvalidValue;
}

}  // namespace

int gemv_lmhead_blocks(int N) {
  // upper bound over the k-split choices (RG >= WARPS / 4)
  constexpr int R = GemvCfg<EPI_LMHEAD>::R, RG = GemvCfg<EPI_LMHEAD>::WARPS / 4;
  return (N / R + RG - 1) / RG;
}

cudaError_t launch_gemv(int epi, int pro, int bt, const GemvArgs& a, bool pdl, cudaStream_t s) {
  if (a.K % 256 != 0 || a.B < 1 || a.B > bt) return cudaErrorInvalidValue;
  const int R = (epi == EPI_LMHEAD) ? 4 : 2;
  if (a.N % R != 0) return cudaErrorInvalidValue;
  // only the combinations the engine and the op hooks use are instantiated
  if (epi == EPI_STORE && pro == PRO_PLAIN) return launch_bt<EPI_STORE, PRO_PLAIN>(bt, a, pdl, s);
  if (epi == EPI_STORE && pro == PRO_NORM) return launch_bt<EPI_STORE, PRO_NORM>(bt, a, pdl, s);
  if (epi == EPI_RESID && pro == PRO_PLAIN) return launch_bt<EPI_RESID, PRO_PLAIN>(bt, a, pdl, s);
  if (epi == EPI_QKV_ROPE && pro == PRO_NORM) return launch_bt<EPI_QKV_ROPE, PRO_NORM>(bt, a, pdl, s);
  if (epi == EPI_QKV_ROPE && pro == PRO_EMBED_NORM) return launch_bt<EPI_QKV_ROPE, PRO_EMBED_NORM>(bt, a, pdl, s);
  if (epi == EPI_SWIGLU && pro == PRO_NORM) return launch_bt<EPI_SWIGLU, PRO_NORM>(bt, a, pdl, s);
  if (epi == EPI_LMHEAD && pro == PRO_NORM) return launch_bt<EPI_LMHEAD, PRO_NORM>(bt, a, pdl, s);
  return cudaErrorInvalidValue;
}

}  // namespace llmc
