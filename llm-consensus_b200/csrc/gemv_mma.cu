// gemv_mma.cu -- the decode-step matrix product for batches of 2..8 sequences: Y[b][n] = sum_k W[n][k] x[b][k].
//
// Still HBM-bound -- W is streamed once per step whatever the batch -- but eight
// activation vectors per weight element are too much FMA and shared-memory traffic for the CUDA cores
// (measured: the FMA kernel of gemv.cu takes 1.59 ms per step at batch 4 against 0.86 at batch 1).  The
// multiply-accumulate therefore runs on the tensor cores with the batch as the N = 8 dimension of
// mma.sync.m16n8k8.tf32:
//   * bf16 weights are exactly representable in tf32 (a 16-bit shift), and every activation is fed as a
//     tf32 pair x = hi + lo (two MMAs), so the products carry fp32-class precision -- the parity budget
//     does not pay for batching;
//   * one block = one 16-row tile (rows chosen so that an accumulator lane holds a RoPE pair / a gate-up
//     pair), its 8 warps split K; a lane loads 16 bytes (8 consecutive k) of one row, and the k order
//     inside a 32-wide block is permuted identically for W and x so that those 8 values are the operands
//     of four consecutive MMAs -- no shuffles, fully used sectors;
//   * activations are staged per 1024-wide K chunk in shared memory, already split into hi/lo, row stride
//     padded so that fragment reads are conflict-free; weight loads run one chunk ahead in registers.
// Fused prologues / epilogues are those of gemv.cu (RMSNorm with folded gain, embedding gather; residual
// add, RoPE + paged-KV append, SwiGLU, greedy argmax + step finalise).
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include "common.cuh"
#include "kernels.h"

namespace llmc {

namespace {

constexpr int XW = 8;            // warps per block (k-slices of one row tile)
constexpr int NB = 8;            // batch columns of the MMA
constexpr int KC = 1024;         // K chunk staged in shared memory
constexpr int XS = KC + 4;       // padded row stride (floats) of the staged activations
constexpr int KB_PER_WARP = KC / 32 / XW;  // 32-wide k blocks per warp per chunk (= 4 -> 8 loads per lane)

__device__ __forceinline__ uint32_t to_tf32(float f) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(f));
  return r;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// rows of tile `tile`: lo[i], hi[i] for i = 0..7 (accumulator rows g and g + 8)
template <int EPI>
__device__ __forceinline__ void tile_rows(const GemvArgs& a, int tile, int i, int& lo, int& hi) {
  if (EPI == EPI_QKV_ROPE) {
    const int hd2 = a.head_dim >> 1, per = hd2 >> 3;
    const int h = tile / per, blk = tile - h * per;
    lo = h * a.head_dim + blk * 8 + i;
    hi = lo + hd2;
  } else if (EPI == EPI_SWIGLU) {
    lo = gu_row_gate(tile * 8 + i);
    hi = gu_row_up(tile * 8 + i);
  } else {
    lo = tile * 16 + i;
    hi = lo + 8;
  }
}

template <int EPI, int PRO>
__global__ void __launch_bounds__(XW * 32, 2) gemv_mma_kernel(const GemvArgs a) {
  extern __shared__ __align__(16) float sm[];
  float* xhi = sm;                  // [NB][XS] tf32 bit patterns
  float* xlo = sm + NB * XS;
  float* part = sm + 2 * NB * XS;   // [XW][32][4] k-slice partial accumulators
  __shared__ float s_ss[XW][NB];
  __shared__ float s_val[XW][NB];
  __shared__ int s_idx[XW][NB];
  __shared__ bool s_last;

  pdl_launch_dependents();

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int K = a.K, nchunks = (K + KC - 1) / KC;
  const int tile = blockIdx.x;
  int row_lo, row_hi;
  tile_rows<EPI>(a, tile, g, row_lo, row_hi);
  const bool lo_ok = row_lo < a.N, hi_ok = row_hi < a.N;
  const uint16_t* wlo = a.W + (size_t)(lo_ok ? row_lo : 0) * K + t * 8;
  const uint16_t* whi = a.W + (size_t)(hi_ok ? row_hi : 0) * K + t * 8;

  // weight registers of one chunk: this warp's KB_PER_WARP k-blocks, rows lo and hi
  uint4 wa[KB_PER_WARP][2], wb[KB_PER_WARP][2];
  auto load_chunk = [&](uint4(&w)[KB_PER_WARP][2], int c) {
#pragma unroll
    for (int j = 0; j < KB_PER_WARP; ++j) {
      const int k0 = c * KC + (warp * KB_PER_WARP + j) * 32;
      const bool in = c < nchunks && k0 < K;
      w[j][0] = in ? ldg_stream(wlo + k0) : make_uint4(0, 0, 0, 0);
      w[j][1] = in ? ldg_stream(whi + k0) : make_uint4(0, 0, 0, 0);
    }
  };
  load_chunk(wa, 0);   // weights do not depend on the previous kernel
  load_chunk(wb, 1);

  pdl_wait();

  // epilogue metadata for the two batch columns of this lane, fetched early
  int e_pos[2] = {0, 0}, e_page[2] = {0, 0};
  float e_cos[2] = {1.f, 1.f}, e_sin[2] = {0.f, 0.f};
  if (EPI == EPI_QKV_ROPE && warp == 0) {
    const int hd2 = a.head_dim >> 1;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int b = 2 * t + j;
      if (b < a.B) {
        const int slot = a.slots[b];
        e_pos[j] = a.seq_len[slot];
        e_page[j] = a.block_table[(size_t)slot * a.bt_stride + e_pos[j] / a.page_size];
        const int i = (row_lo % a.head_dim);
        e_cos[j] = a.rope_cos[(size_t)e_pos[j] * hd2 + i];
        e_sin[j] = a.rope_sin[(size_t)e_pos[j] * hd2 + i];
      }
    }
  }

  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  float ss_local[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) ss_local[b] = 0.f;

  auto stage = [&](int c) {
    // x chunk c -> shared memory as tf32 hi / lo, gain folded; thread = 4 consecutive k of every batch row
    for (int k4 = tid * 4; k4 < KC; k4 += XW * 32 * 4) {
      const int k = c * KC + k4;
      float4 gv = make_float4(1.f, 1.f, 1.f, 1.f);
      if (PRO != PRO_PLAIN && k < K) {
        const uint2 gb = *reinterpret_cast<const uint2*>(a.gain + k);
        gv = make_float4(bf16lo(gb.x), bf16hi(gb.x), bf16lo(gb.y), bf16hi(gb.y));
      }
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (b < a.B && k < K) {
          if (PRO == PRO_EMBED_NORM) {
            const int tok = a.next_token[a.slots[b]];
            const uint2 e = *reinterpret_cast<const uint2*>(a.emb + (size_t)tok * K + k);
            v = make_float4(bf16lo(e.x), bf16hi(e.x), bf16lo(e.y), bf16hi(e.y));
            if (blockIdx.x == 0) *reinterpret_cast<float4*>(a.x_out + (size_t)b * K + k) = v;
          } else {
            v = *reinterpret_cast<const float4*>(a.x + (size_t)b * a.x_stride + k);
          }
        }
        if (PRO != PRO_PLAIN) ss_local[b] += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        v.x *= gv.x; v.y *= gv.y; v.z *= gv.z; v.w *= gv.w;
        float4 h, l;
        h.x = __uint_as_float(to_tf32(v.x)); l.x = __uint_as_float(to_tf32(v.x - h.x));
        h.y = __uint_as_float(to_tf32(v.y)); l.y = __uint_as_float(to_tf32(v.y - h.y));
        h.z = __uint_as_float(to_tf32(v.z)); l.z = __uint_as_float(to_tf32(v.z - h.z));
        h.w = __uint_as_float(to_tf32(v.w)); l.w = __uint_as_float(to_tf32(v.w - h.w));
        *reinterpret_cast<float4*>(xhi + b * XS + k4) = h;
        *reinterpret_cast<float4*>(xlo + b * XS + k4) = l;
      }
    }
  };

  auto compute = [&](uint4(&w)[KB_PER_WARP][2], int c) {
#pragma unroll
    for (int j = 0; j < KB_PER_WARP; ++j) {
      const int kk = (warp * KB_PER_WARP + j) * 32 + t * 8;      // this lane's 8 consecutive k inside the chunk
      // the k order inside the 32-block is permuted the same way for W and x: MMA step s uses k = 2s, 2s+1
      const float4 h0 = *reinterpret_cast<const float4*>(xhi + g * XS + kk);
      const float4 h1 = *reinterpret_cast<const float4*>(xhi + g * XS + kk + 4);
      const float4 l0 = *reinterpret_cast<const float4*>(xlo + g * XS + kk);
      const float4 l1 = *reinterpret_cast<const float4*>(xlo + g * XS + kk + 4);
      const uint32_t wl[4] = {w[j][0].x, w[j][0].y, w[j][0].z, w[j][0].w};
      const uint32_t wh[4] = {w[j][1].x, w[j][1].y, w[j][1].z, w[j][1].w};
      const float xh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
      const float xl[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const uint32_t a0 = wl[s] << 16, a2 = wl[s] & 0xffff0000u;   // bf16 -> tf32 is exact
        const uint32_t a1 = wh[s] << 16, a3 = wh[s] & 0xffff0000u;
        mma_tf32(acc, a0, a1, a2, a3, __float_as_uint(xh[2 * s]), __float_as_uint(xh[2 * s + 1]));
        mma_tf32(acc, a0, a1, a2, a3, __float_as_uint(xl[2 * s]), __float_as_uint(xl[2 * s + 1]));
      }
    }
    (void)c;
  };

  for (int c = 0; c < nchunks; c += 2) {
    stage(c);
    __syncthreads();
    compute(wa, c);
    load_chunk(wa, c + 2);
    __syncthreads();
    if (c + 1 < nchunks) {
      stage(c + 1);
      __syncthreads();
      compute(wb, c + 1);
      load_chunk(wb, c + 3);
      __syncthreads();
    }
  }

  // ---- combine the 8 k-slices (and the per-batch sums of squares)
  *reinterpret_cast<float4*>(part + ((size_t)warp * 32 + lane) * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  if (PRO != PRO_PLAIN) {
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const float s = warp_sum(ss_local[b]);
      if (lane == 0) s_ss[warp][b] = s;
    }
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int w = 1; w < XW; ++w) {
      const float4 p = *reinterpret_cast<const float4*>(part + ((size_t)w * 32 + lane) * 4);
      acc[0] += p.x; acc[1] += p.y; acc[2] += p.z; acc[3] += p.w;
    }
    // this lane: rows (row_lo, row_hi) x batches (2t, 2t+1): acc = {lo/b0, lo/b1, hi/b0, hi/b1}
    float rs[2] = {1.f, 1.f};
    if (PRO != PRO_PLAIN) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < XW; ++w) s += s_ss[w][2 * t + j];
        rs[j] = rsqrtf(s / (float)K + a.eps);
      }
    }
    float v_lo[2] = {acc[0] * rs[0], acc[1] * rs[1]}, v_hi[2] = {acc[2] * rs[0], acc[3] * rs[1]};
    float best[2] = {-INFINITY, -INFINITY};
    int bidx[2] = {0x7fffffff, 0x7fffffff};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int b = 2 * t + j;
      if (b >= a.B) continue;
      if (EPI == EPI_STORE || EPI == EPI_RESID) {
        if (lo_ok) {
          float* p = a.y + (size_t)b * a.y_stride + row_lo;
          *p = (EPI == EPI_RESID) ? *p + v_lo[j] : v_lo[j];
        }
        if (hi_ok) {
          float* p = a.y + (size_t)b * a.y_stride + row_hi;
          *p = (EPI == EPI_RESID) ? *p + v_hi[j] : v_hi[j];
        }
      } else if (EPI == EPI_SWIGLU) {
        a.y[(size_t)b * a.y_stride + tile * 8 + g] = v_lo[j] / (1.0f + __expf(-v_lo[j])) * v_hi[j];
      } else if (EPI == EPI_QKV_ROPE) {
        const int d = a.head_dim, hd2 = d >> 1;
        const int h = row_lo / d, i = row_lo - h * d;
        float x0 = v_lo[j], x1 = v_hi[j];
        if (h < a.q_heads + a.kv_heads) {
          const float o0 = x0 * e_cos[j] - x1 * e_sin[j], o1 = x1 * e_cos[j] + x0 * e_sin[j];
          x0 = o0; x1 = o1;
        }
        if (h < a.q_heads) {
          float* qp = a.y + (size_t)b * a.y_stride + h * d + i;
          qp[0] = x0;
          qp[hd2] = x1;
        } else {
          const bool isk = h < a.q_heads + a.kv_heads;
          const int kvh = isk ? h - a.q_heads : h - a.q_heads - a.kv_heads;
          const size_t eo = (((size_t)e_page[j] * a.kv_heads + kvh) * a.page_size + (e_pos[j] % a.page_size)) * d + i;
          if (a.kv_fp32) {
            float* dst = reinterpret_cast<float*>(isk ? a.kpool : a.vpool) + eo;
            dst[0] = x0;
            dst[hd2] = x1;
          } else {
            uint16_t* dst = reinterpret_cast<uint16_t*>(isk ? a.kpool : a.vpool) + eo;
            dst[0] = f32_to_bf16_bits(x0);
            dst[hd2] = f32_to_bf16_bits(x1);
          }
        }
      } else if (EPI == EPI_LMHEAD) {
        if (lo_ok) {
          if (a.y) a.y[(size_t)b * a.y_stride + row_lo] = v_lo[j];
          best[j] = v_lo[j];
          bidx[j] = row_lo;
        }
        if (hi_ok) {
          if (a.y) a.y[(size_t)b * a.y_stride + row_hi] = v_hi[j];
          if (v_hi[j] > best[j]) { best[j] = v_hi[j]; bidx[j] = row_hi; }
        }
      }
    }
    if (EPI == EPI_LMHEAD) {
      // per batch column: best over the 16 rows of the tile = over the 8 lanes sharing t
#pragma unroll
      for (int j = 0; j < 2; ++j) {
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, best[j], o);
          const int oi = __shfl_xor_sync(0xffffffffu, bidx[j], o);
          if (ov > best[j] || (ov == best[j] && oi < bidx[j])) { best[j] = ov; bidx[j] = oi; }
        }
        if (g == 0 && 2 * t + j < a.B) {
          a.part_val[(size_t)(2 * t + j) * gridDim.x + blockIdx.x] = best[j];
          a.part_idx[(size_t)(2 * t + j) * gridDim.x + blockIdx.x] = bidx[j];
        }
      }
    }
  }
  if (EPI != EPI_LMHEAD) return;

  // ---- LM head: the last block reduces the per-block candidates and finalises the step (as in gemv.cu)
  const int nblk = gridDim.x;
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(a.counter, 1u) == (unsigned)(nblk - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int b = 0; b < a.B; ++b) {
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int j = tid; j < nblk; j += XW * 32) {
      const float v = __ldcg(a.part_val + (size_t)b * nblk + j);
      const int ix = __ldcg(a.part_idx + (size_t)b * nblk + j);
      if (v > best || (v == best && ix < bi)) { best = v; bi = ix; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    __syncthreads();
    if (lane == 0) { s_val[warp][0] = best; s_idx[warp][0] = bi; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < XW; ++w)
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

template <int EPI>
int num_tiles(const GemvArgs& a) {
  if (EPI == EPI_SWIGLU) return a.N / 16;   // 8 gate + 8 up rows per tile
  if (EPI == EPI_QKV_ROPE) return a.N / 16;
  return (a.N + 15) / 16;
}

template <int EPI, int PRO>
cudaError_t launch_mma(const GemvArgs& a, bool pdl, cudaStream_t s) {
  const size_t smem = ((size_t)2 * NB * XS + (size_t)XW * 32 * 4) * sizeof(float);
  auto kern = gemv_mma_kernel<EPI, PRO>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_tiles<EPI>(a));
  cfg.blockDim = dim3(XW * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, a);
}

}  // namespace

int gemv_mma_lmhead_blocks(int N) { return (N + 15) / 16; }

cudaError_t launch_gemv_mma(int epi, int pro, const GemvArgs& a, bool pdl, cudaStream_t s) {
  if (a.K % 32 != 0 || a.B < 1 || a.B > NB) return cudaErrorInvalidValue;
  if ((epi == EPI_QKV_ROPE || epi == EPI_SWIGLU) && a.N % 16 != 0) return cudaErrorInvalidValue;
  if (epi == EPI_QKV_ROPE && (a.head_dim / 2) % 8 != 0) return cudaErrorInvalidValue;
  if (epi == EPI_STORE && pro == PRO_PLAIN) return launch_mma<EPI_STORE, PRO_PLAIN>(a, pdl, s);
  if (epi == EPI_STORE && pro == PRO_NORM) return launch_mma<EPI_STORE, PRO_NORM>(a, pdl, s);
  if (epi == EPI_RESID && pro == PRO_PLAIN) return launch_mma<EPI_RESID, PRO_PLAIN>(a, pdl, s);
  if (epi == EPI_QKV_ROPE && pro == PRO_NORM) return launch_mma<EPI_QKV_ROPE, PRO_NORM>(a, pdl, s);
  if (epi == EPI_QKV_ROPE && pro == PRO_EMBED_NORM) return launch_mma<EPI_QKV_ROPE, PRO_EMBED_NORM>(a, pdl, s);
  if (epi == EPI_SWIGLU && pro == PRO_NORM) return launch_mma<EPI_SWIGLU, PRO_NORM>(a, pdl, s);
  if (epi == EPI_LMHEAD && pro == PRO_NORM) return launch_mma<EPI_LMHEAD, PRO_NORM>(a, pdl, s);
  return cudaErrorInvalidValue;
}

}  // namespace llmc
