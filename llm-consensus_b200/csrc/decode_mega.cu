// decode_mega.cu -- the batch-1 decode step as ONE persistent cooperative kernel.
//
// A decode step of the 1B shape is 81 dependent phases (5 per layer + LM head) of 8-67 MB each, i.e.
// 1-10 us of HBM time per phase.  Launched as 81 kernels, each phase pays a launch, a cold start of
// its weight stream and a drain; measured, that is ~2.3x the bytes/bandwidth time (profiles/README.md).
// Here the whole step -- and all `nsteps` steps of a llmc_decode call -- run inside one kernel:
//   * grid = one block of 16 warps per SM (cooperative launch, all blocks co-resident); a phase's
//     row pairs are dealt round-robin to the 148 x 16 warps, each warp streaming its rows with
//     16-byte coalesced L1::no_allocate loads, two register batches of 8 loads in flight, continuing
//     across its tasks;
//   * phases are separated by a grid barrier (one atomic counter, ~1 us); BEFORE it arrives at the
//     barrier every warp requests the first two weight batches of its first task of the NEXT phase
//     (weights never depend on a barrier), so the stream is already in flight while the barrier and
//     the re-staging of the activation vector happen;
//   * activations cross between SMs through L2 (ld.global.cg; L1 is not coherent inside a kernel);
//   * the attention phase gives (kv head, split) to blocks and 32-position tiles to their warps
//     (attn_tile.cuh); tile states merge in shared memory, and the last block to finish a kv head (one
//     atomic counter per head, no atomics on data) merges the <= 8 split records of its heads;
//   * the LM-head phase ends in a per-block argmax candidate; after the barrier every block reduces
//     the 148 candidates itself, so the next step starts with no further exchange.
// HBM-bound; algorithmic bytes per step as for the multi-kernel path.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the
// boundary; the providers behind it are remote HTTP calls).
#include <cooperative_groups.h>

#include "attn_tile.cuh"
#include "common.cuh"
#include "kernels.h"

namespace llmc {

namespace {

constexpr int MW = 16;      // warps per block
constexpr int MUK = 4;      // k-iterations per batch: 2 rows x 4 = 8 loads per lane per batch
constexpr int MEGA_MAX_SPLIT = 8;  // blocks per kv head in the attention phase
enum { MODE_QKV = 0, MODE_RESID = 1, MODE_SWIGLU = 2, MODE_LMHEAD = 3 };

struct PhaseDesc {
  const uint16_t* W;
  int ntasks;  // row pairs
  int K;
  int mode;
  int hd2;     // MODE_QKV: head_dim / 2
};

__device__ __forceinline__ void task_rows(const PhaseDesc& p, int g, int& r0, int& r1) {
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

// The warp's work in a phase is a flat sequence of batches f = 0, 1, ... over (task, k-range):
// task = gwarp + (f / nb) * nwarps, k-iterations [(f % nb) * MUK, +MUK).
struct WarpStream {
  int gwarp, nwarps, lane;
};

__device__ __forceinline__ void load_flat(const PhaseDesc& p, const WarpStream& ws, int f, uint4 (&w)[2][MUK]) {
  const int nit = p.K >> 8;
  const int span = MUK;  // k-iterations covered by one batch
  const int nb = (nit + span - 1) / span;
  const int task = ws.gwarp + (f / nb) * ws.nwarps;
  const int it0 = (f % nb) * span;
  if (task < p.ntasks) {
    int r0, r1;
    task_rows(p, task, r0, r1);
    const uint16_t* p0 = p.W + (size_t)r0 * p.K + ws.lane * 8;
    const uint16_t* p1 = p.W + (size_t)r1 * p.K + ws.lane * 8;
#pragma unroll
    for (int u = 0; u < MUK; ++u) {
      const int it = it0 + u;
      w[0][u] = it < nit ? ldg_stream(p0 + (size_t)it * 256) : make_uint4(0, 0, 0, 0);
      w[1][u] = it < nit ? ldg_stream(p1 + (size_t)it * 256) : make_uint4(0, 0, 0, 0);
    }
  } else {
#pragma unroll
    for (int u = 0; u < MUK; ++u) {
      w[0][u] = make_uint4(0, 0, 0, 0);
      w[1][u] = make_uint4(0, 0, 0, 0);
    }
  }
}

__device__ __forceinline__ int xs_index(int k) {  // same permuted layout as gemv.cu
  const int c = k >> 8, w = k & 255;
  const int lane = w >> 3, j = w & 7;
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

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// trace (optional, llmc_config.flags & LLMC_FLAG_TRACE): block 0 records, per barrier, the time it
// arrived and the time the barrier opened: [2 * index], [2 * index + 1] in ns
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& target, unsigned long long* trace, int& tix) {
  __syncthreads();
  if (threadIdx.x == 0) {
    if (trace && blockIdx.x == 0) trace[2 * tix] = gtime();
    // release-add / acquire-poll at gpu scope: the block's writes (ordered before this thread by the
    // barrier above) become visible to every block that observes the count, without a full membar
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
    unsigned v;   // relaxed spin, then one acquire load (cheaper than spinning on acquire loads; see decode_ring.cu)
    do {
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
    } while (v < target);
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
    if (trace && blockIdx.x == 0) trace[2 * tix + 1] = gtime();
  }
  __syncthreads();
  target += gridDim.x;
  ++tix;
}

// Stage an activation vector into shared memory (block-wide).  emb != NULL: the row of a bf16
// embedding table; otherwise fp32 from L2.  gain != NULL: fold the RMSNorm gain and return 1/rms.
__device__ __forceinline__ float stage_x(const float* src, const uint16_t* emb_row, const uint16_t* gain, int K,
                                         float eps, float* xs, float* red, float* x_out) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float ss = 0.f;
  // one 16-byte load per thread per 2048 elements: the prefetched weight batches are live here, so this
  // stays register-light on purpose (wider batching or cp.async staging both measured slower)
  for (int k = tid * 4; k < K; k += MW * 32 * 4) {
    float4 v;
    if (emb_row) {
      const uint2 e = *reinterpret_cast<const uint2*>(emb_row + k);
      v = make_float4(bf16lo(e.x), bf16hi(e.x), bf16lo(e.y), bf16hi(e.y));
      if (x_out) *reinterpret_cast<float4*>(x_out + k) = v;
    } else {
      v = __ldcg(reinterpret_cast<const float4*>(src + k));
    }
    if (gain) {
      ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
      const uint2 gb = *reinterpret_cast<const uint2*>(gain + k);
      v.x *= bf16lo(gb.x); v.y *= bf16hi(gb.x); v.z *= bf16lo(gb.y); v.w *= bf16hi(gb.y);
    }
    *reinterpret_cast<float4*>(xs + xs_index(k)) = v;
  }
  float rs = 1.f;
  if (gain) {
    ss = warp_sum(ss);
    if (lane == 0) red[warp] = ss;
  }
  __syncthreads();
  if (gain) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < MW; ++w) t += red[w];
    rs = rsqrtf(t / (float)K + eps);
  }
  return rs;
}

template <int D, int G, typename KVT>
__global__ void __launch_bounds__(MW * 32, 1) decode_mega_kernel(const MegaArgs a) {
  constexpr int PS = D + 2;
  using LM = LaneMap<D, G, KVT>;
  extern __shared__ __align__(16) float sm[];
  float* xs = sm;                 // staged activation vector; the attention phase reuses it
  float* red = sm + a.smem_x;     // [MW]
  __shared__ float s_val[MW];
  __shared__ int s_idx[MW];
  __shared__ int s_tok;
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  WarpStream ws;
  ws.lane = lane;
  ws.nwarps = gridDim.x * MW;
  ws.gwarp = warp * gridDim.x + blockIdx.x;  // consecutive tasks land on different SMs
  unsigned bar_target = gridDim.x;
  int tix = 0;

  const int H = a.H, QD = a.QD, I = a.I, V = a.V, hd2 = D / 2;
  const int slot = a.slot;
  const int pos0 = a.seq_len[slot];
  int tok = a.next_token[slot];
  const int* bt = a.block_table + (size_t)slot * a.bt_stride;
  const float scale = rsqrtf((float)D);

  uint4 wcur[2][MUK], wnxt[2][MUK];

  // one GEMV phase over the staged vector xs; wcur/wnxt already hold flat batches 0 and 1.
  // Ping-pong: a batch's registers are re-requested right after it has been consumed and are not
  // touched again until the OTHER batch has been consumed, so two batches (16 loads per lane) stay in
  // flight.  (Copying wnxt into wcur instead would make the copy wait for the loads in flight.)
  auto run_phase = [&](const PhaseDesc& p, float rs, auto&& epilogue) {
    const int nit = p.K >> 8;
    const int span = MUK;
    const int nb = (nit + span - 1) / span;
    const int my_tasks = ws.gwarp < p.ntasks ? (p.ntasks - ws.gwarp + ws.nwarps - 1) / ws.nwarps : 0;
    const int nflat = my_tasks * nb;
    float acc0 = 0.f, acc1 = 0.f;
    auto consume = [&](uint4 (&w)[2][MUK], int f) {
      const int it0 = (f % nb) * span;
#pragma unroll
      for (int u = 0; u < MUK; ++u) {
        const int it = it0 + u;
        if (it < nit) {
          const float* xp = xs + (it << 8) + (lane << 2);
          const float4 x0 = *reinterpret_cast<const float4*>(xp);
          const float4 x1 = *reinterpret_cast<const float4*>(xp + 128);
          acc0 += dot8(w[0][u], x0, x1);
          acc1 += dot8(w[1][u], x0, x1);
        }

      }
      load_flat(p, ws, f + 2, w);  // refill the registers just consumed
      if ((f % nb) == nb - 1) {    // task complete
        const int task = ws.gwarp + (f / nb) * ws.nwarps;
        acc0 = warp_sum(acc0) * rs;
        acc1 = warp_sum(acc1) * rs;
        epilogue(task, acc0, acc1);
        acc0 = 0.f;
        acc1 = 0.f;
      }
    };
    for (int f = 0; f < nflat; f += 2) {
      consume(wcur, f);
      if (f + 1 < nflat) consume(wnxt, f + 1);
    }
  };
  auto prefetch = [&](const PhaseDesc& p) {
    load_flat(p, ws, 0, wcur);
    load_flat(p, ws, 1, wnxt);
  };
  auto resid_epilogue = [&](int task, float v0, float v1) {
    if (lane == 0) {
      float2* px = reinterpret_cast<float2*>(a.x + 2 * task);
      float2 old = __ldcg(px);
      old.x += v0;
      old.y += v1;
      *px = old;
    }
  };

  PhaseDesc pd;
  pd.W = a.layers[0].wqkv; pd.ntasks = (QD + 2 * a.KD) / 2; pd.K = H; pd.mode = MODE_QKV; pd.hd2 = hd2;
  prefetch(pd);

  for (int step = 0; step < a.nsteps; ++step) {
    const int pos = pos0 + step;
    const int ctx = pos + 1;
    const int ntiles = (ctx + TP - 1) / TP;
    const int page_new = bt[pos / a.page_size];
    const int off_new = pos % a.page_size;

    for (int l = 0; l < a.L; ++l) {
      const MegaLayer& ly = a.layers[l];
      KVT* kpool = reinterpret_cast<KVT*>(ly.kp);
      KVT* vpool = reinterpret_cast<KVT*>(ly.vp);

      // ---------------- P1: RMSNorm -> QKV -> RoPE -> KV append
      {
        const float rs = stage_x(a.x, l == 0 ? a.emb + (size_t)tok * H : nullptr, ly.g1, H, a.eps, xs, red,
                                 (l == 0 && blockIdx.x == 0) ? a.x : nullptr);
        pd.W = ly.wqkv; pd.ntasks = (QD + 2 * a.KD) / 2; pd.K = H; pd.mode = MODE_QKV; pd.hd2 = hd2;
        run_phase(pd, rs, [&](int g, float v0, float v1) {
          if (lane == 0) {
            const int h = g / hd2, i = g - h * hd2;
            if (h < a.q_heads + a.kv_heads) {
              const float c = a.rope_cos[(size_t)pos * hd2 + i], s = a.rope_sin[(size_t)pos * hd2 + i];
              const float o0 = v0 * c - v1 * s, o1 = v1 * c + v0 * s;
              v0 = o0; v1 = o1;
            }
            if (h < a.q_heads) {
              a.q[h * D + i] = v0;
              a.q[h * D + i + hd2] = v1;
            } else {
              const bool isk = h < a.q_heads + a.kv_heads;
              const int kvh = isk ? h - a.q_heads : h - a.q_heads - a.kv_heads;
              KVT* dst = (isk ? kpool : vpool) + (((size_t)page_new * a.kv_heads + kvh) * a.page_size + off_new) * D + i;
              if constexpr (sizeof(KVT) == 4) {
                dst[0] = v0;
                dst[hd2] = v1;
              } else {
                dst[0] = f32_to_bf16_bits(v0);
                dst[hd2] = f32_to_bf16_bits(v1);
              }
            }
          }
        });
        // K/V tiles of this step's attention were evicted from L2 by the weight stream: request them now, so the two
        // dependent tile loads of the attention phase cost an L2 trip instead of an HBM trip (as in decode_ring.cu)
        {
          const int nsp_pf = min(MEGA_MAX_SPLIT, max(1, (ntiles + a.attn_tiles_per_block - 1) / a.attn_tiles_per_block));
          if ((int)blockIdx.x < a.kv_heads * nsp_pf) {
            const int kvh = blockIdx.x % a.kv_heads, split = blockIdx.x / a.kv_heads;
            constexpr int TILE_LINES = TP * D * (int)sizeof(KVT) / 128;
            for (int tile = split + nsp_pf * warp; tile < ntiles; tile += nsp_pf * MW) {
              const int p0 = tile * TP;
              const int page = bt[p0 / a.page_size];
              const size_t base = (((size_t)page * a.kv_heads + kvh) * a.page_size + (p0 % a.page_size)) * D;
              const char* kb = reinterpret_cast<const char*>(kpool + base);
              const char* vb = reinterpret_cast<const char*>(vpool + base);
              for (int i = lane; i < TILE_LINES; i += 32) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(kb + (size_t)i * 128));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(vb + (size_t)i * 128));
              }
            }
          }
        }
        grid_barrier(a.barrier, bar_target, a.trace, tix);
      }

      // ---------------- P2: attention.  Block b < kv_heads * nsp takes (kv head, split): its warps take one
      // 32-position tile each (tiles split, split + nsp, ...), their softmax states meet in shared memory and
      // one record per (kv head, split, q head) goes to the scratch.  nsp is chosen so that a block has
      // about 4 tiles: the work spreads over tens of SMs, and the cross-block merge (P3) has few terms.
      const int nsp = min(MEGA_MAX_SPLIT, max(1, (ntiles + a.attn_tiles_per_block - 1) / a.attn_tiles_per_block));
      {
        float* qs = xs;                                            // [G][D]
        float* part = xs + G * D;                                  // [MW][G][PS]
        float* pw = part + (size_t)MW * G * PS + (size_t)warp * G * (TP + 1);
        if ((int)blockIdx.x < a.kv_heads * nsp) {
          const int kvh = blockIdx.x % a.kv_heads, split = blockIdx.x / a.kv_heads;
          for (int i = tid; i < G * D; i += MW * 32) qs[i] = __ldcg(a.q + (size_t)kvh * G * D + i);
          __syncthreads();
          const int c = lane % LM::LPR;
          float m[G], lsum[G], o[G][LM::EPL];
#pragma unroll
          for (int h = 0; h < G; ++h) {
            m[h] = -INFINITY;
            lsum[h] = 0.f;
#pragma unroll
            for (int e = 0; e < LM::EPL; ++e) o[h][e] = 0.f;
          }
          for (int tile = split + nsp * warp; tile < ntiles; tile += nsp * MW) {
            const int p0 = tile * TP;
            const int valid = min(TP, ctx - p0);
            const int page = bt[p0 / a.page_size];
            const size_t base = (((size_t)page * a.kv_heads + kvh) * a.page_size + (p0 % a.page_size)) * D;
            attn_tile_step<D, G, KVT, true>(qs, pw, kpool + base + c * LM::EPV, vpool + base + c * LM::EPV, valid,
                                            scale, m, lsum, o);
          }
          attn_write_record<D, G, KVT>(part + (size_t)warp * G * PS, m, lsum, o);
          __syncthreads();
          if (warp < G) {  // warp h merges head h over the block's warps
            constexpr int DPL = D / 32;
            const int h = warp;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < MW; ++w) M = fmaxf(M, part[((size_t)w * G + h) * PS + D]);
            float L = 0.f, O[DPL];
#pragma unroll
            for (int i = 0; i < DPL; ++i) O[i] = 0.f;
            if (M != -INFINITY) {
#pragma unroll
              for (int w = 0; w < MW; ++w) {
                const float* pp = part + ((size_t)w * G + h) * PS;
                const float mw = pp[D];
                const float wt = (mw == -INFINITY) ? 0.f : __expf(mw - M);
                L = fmaf(pp[D + 1], wt, L);
#pragma unroll
                for (int i = 0; i < DPL; ++i) O[i] = fmaf(pp[lane * DPL + i], wt, O[i]);
              }
            }
            float* rec = a.attn_scratch + (((size_t)kvh * MEGA_MAX_SPLIT + split) * G + h) * PS;
#pragma unroll
            for (int i = 0; i < DPL; ++i) rec[lane * DPL + i] = O[i];
            if (lane == 0) {
              rec[D] = M;
              rec[D + 1] = L;
            }
          }
          // the last block to finish this kv head (one atomic counter per head, no atomics on data) merges
          // its nsp split records: warp h takes head h, all record loads of a lane issued together
          __threadfence();
          __syncthreads();
          if (tid == 0) s_last = (atomicAdd(a.attn_counters + kvh, 1u) == (unsigned)(nsp - 1)) ? 1 : 0;
          __syncthreads();
          if (s_last && warp < G) {
            __threadfence();
            constexpr int DPL = D / 32;
            const int h = warp;
            const float* rb = a.attn_scratch + ((size_t)kvh * MEGA_MAX_SPLIT * G + h) * PS;
            float mv[MEGA_MAX_SPLIT], lv[MEGA_MAX_SPLIT], ov[MEGA_MAX_SPLIT][DPL];
#pragma unroll
            for (int sp = 0; sp < MEGA_MAX_SPLIT; ++sp) {
              const float* r = rb + (size_t)sp * G * PS;
              const bool live = sp < nsp;
              mv[sp] = live ? __ldcg(r + D) : -INFINITY;
              lv[sp] = live ? __ldcg(r + D + 1) : 0.f;
#pragma unroll
              for (int i = 0; i < DPL; ++i) ov[sp][i] = live ? __ldcg(r + lane * DPL + i) : 0.f;
            }
            float M = -INFINITY;
#pragma unroll
            for (int sp = 0; sp < MEGA_MAX_SPLIT; ++sp) M = fmaxf(M, mv[sp]);
            float L = 0.f, O[DPL];
#pragma unroll
            for (int i = 0; i < DPL; ++i) O[i] = 0.f;
#pragma unroll
            for (int sp = 0; sp < MEGA_MAX_SPLIT; ++sp) {
              const float wt = (mv[sp] == -INFINITY) ? 0.f : __expf(mv[sp] - M);
              L = fmaf(lv[sp], wt, L);
#pragma unroll
              for (int i = 0; i < DPL; ++i) O[i] = fmaf(ov[sp][i], wt, O[i]);
            }
            const float inv = 1.0f / L;
#pragma unroll
            for (int i = 0; i < DPL; ++i) a.attn[(size_t)(kvh * G + h) * D + lane * DPL + i] = O[i] * inv;
            if (lane == 0 && h == 0) a.attn_counters[kvh] = 0u;
          }
        }
        // O-proj weights are requested before the barrier (blocks without attention work get here at once
        // and stream the whole O matrix while the attention runs)
        pd.W = ly.wo; pd.ntasks = H / 2; pd.K = QD; pd.mode = MODE_RESID;
        prefetch(pd);
        grid_barrier(a.barrier, bar_target, a.trace, tix);
      }

      // ---------------- P3: x += Wo attn
      {
        stage_x(a.attn, nullptr, nullptr, QD, a.eps, xs, red, nullptr);
        pd.W = ly.wo; pd.ntasks = H / 2; pd.K = QD; pd.mode = MODE_RESID;
        run_phase(pd, 1.f, resid_epilogue);
        pd.W = ly.wgu; pd.ntasks = I; pd.K = H; pd.mode = MODE_SWIGLU;
        prefetch(pd);
        grid_barrier(a.barrier, bar_target, a.trace, tix);
      }

      // ---------------- P4: act = silu(Wg h) * (Wu h), h = RMSNorm(x)
      {
        const float rs = stage_x(a.x, nullptr, ly.g2, H, a.eps, xs, red, nullptr);
        pd.W = ly.wgu; pd.ntasks = I; pd.K = H; pd.mode = MODE_SWIGLU;
        run_phase(pd, rs, [&](int g, float gt, float up) {
          if (lane == 0) a.act[g] = gt / (1.0f + __expf(-gt)) * up;
        });
        pd.W = ly.wd; pd.ntasks = H / 2; pd.K = I; pd.mode = MODE_RESID;
        prefetch(pd);
        grid_barrier(a.barrier, bar_target, a.trace, tix);
      }

      // ---------------- P5: x += Wd act
      {
        stage_x(a.act, nullptr, nullptr, I, a.eps, xs, red, nullptr);
        pd.W = ly.wd; pd.ntasks = H / 2; pd.K = I; pd.mode = MODE_RESID;
        run_phase(pd, 1.f, resid_epilogue);
        if (l + 1 < a.L) {
          pd.W = a.layers[l + 1].wqkv; pd.ntasks = (QD + 2 * a.KD) / 2; pd.K = H; pd.mode = MODE_QKV; pd.hd2 = hd2;
        } else {
          pd.W = a.lmhead; pd.ntasks = V / 2; pd.K = H; pd.mode = MODE_LMHEAD;
        }
        prefetch(pd);
        grid_barrier(a.barrier, bar_target, a.trace, tix);
      }
    }

    // ---------------- LM head + greedy argmax (lowest index wins ties)
    {
      const float rs = stage_x(a.x, nullptr, a.gf, H, a.eps, xs, red, nullptr);
      pd.W = a.lmhead; pd.ntasks = V / 2; pd.K = H; pd.mode = MODE_LMHEAD;
      float best = -INFINITY;
      int bi = 0x7fffffff;
      const bool store_logits = a.logits != nullptr && step == a.nsteps - 1;
      run_phase(pd, rs, [&](int g, float v0, float v1) {
        if (store_logits && lane == 0) *reinterpret_cast<float2*>(a.logits + 2 * g) = make_float2(v0, v1);
        if (v0 > best || (v0 == best && 2 * g < bi)) { best = v0; bi = 2 * g; }
        if (v1 > best || (v1 == best && 2 * g + 1 < bi)) { best = v1; bi = 2 * g + 1; }
      });
      if (lane == 0) {
        s_val[warp] = best;
        s_idx[warp] = bi;
      }
      __syncthreads();
      if (tid == 0) {
#pragma unroll
        for (int w = 1; w < MW; ++w)
          if (s_val[w] > best || (s_val[w] == best && s_idx[w] < bi)) { best = s_val[w]; bi = s_idx[w]; }
        a.part_val[blockIdx.x] = best;
        a.part_idx[blockIdx.x] = bi;
      }
      // next step's first phase (layer 0 QKV) can already start streaming
      pd.W = a.layers[0].wqkv; pd.ntasks = (QD + 2 * a.KD) / 2; pd.K = H; pd.mode = MODE_QKV; pd.hd2 = hd2;
      if (step + 1 < a.nsteps) prefetch(pd);
      grid_barrier(a.barrier, bar_target, a.trace, tix);
      // every block reduces the per-block candidates itself: no second exchange
      if (warp == 0) {
        float bv = -INFINITY;
        int bx = 0x7fffffff;
        for (int j = lane; j < (int)gridDim.x; j += 32) {
          const float v = __ldcg(a.part_val + j);
          const int ix = __ldcg(a.part_idx + j);
          if (v > bv || (v == bv && ix < bx)) { bv = v; bx = ix; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bx, o);
          if (ov > bv || (ov == bv && oi < bx)) { bv = ov; bx = oi; }
        }
        if (lane == 0) {
          s_tok = bx;
          if (blockIdx.x == 0) a.out_tokens[step] = bx;
        }
      }
      __syncthreads();
      tok = s_tok;
      // part_val/part_idx are rewritten only after the next step's 5*L barriers: no hazard
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    a.next_token[slot] = tok;
    a.seq_len[slot] = pos0 + a.nsteps;
  }
}

template <int D, int G, typename KVT>
cudaError_t launch_mega_t(const MegaArgs& a, size_t smem, int grid, cudaStream_t s) {
  auto kern = decode_mega_kernel<D, G, KVT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  MegaArgs copy = a;
  void* params[] = {&copy};
  return cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3(MW * 32), params, smem, s);
}

template <int D, typename KVT>
cudaError_t launch_mega_g(int G, const MegaArgs& a, size_t smem, int grid, cudaStream_t s) {
  switch (G) {
    case 1: return launch_mega_t<D, 1, KVT>(a, smem, grid, s);
    case 2: return launch_mega_t<D, 2, KVT>(a, smem, grid, s);
    case 4: return launch_mega_t<D, 4, KVT>(a, smem, grid, s);
    case 8: return launch_mega_t<D, 8, KVT>(a, smem, grid, s);
  }
  return cudaErrorInvalidValue;
}

template <typename KVT>
cudaError_t launch_mega_d(const MegaArgs& a, size_t smem, int grid, cudaStream_t s) {
  const int G = a.q_heads / a.kv_heads;
  switch (a.head_dim) {
    case 32: return launch_mega_g<32, KVT>(G, a, smem, grid, s);
    case 64: return launch_mega_g<64, KVT>(G, a, smem, grid, s);
    case 128: return launch_mega_g<128, KVT>(G, a, smem, grid, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

size_t decode_mega_smem_floats(int H, int QD, int I, int G, int D) {
  size_t kx = (size_t)(H > QD ? H : QD);
  if ((size_t)I > kx) kx = I;
  const size_t att = (size_t)G * D + (size_t)MW * G * (D + 2) + (size_t)MW * G * (TP + 1);
  return kx > att ? kx : att;
}

cudaError_t launch_decode_mega(const MegaArgs& a, int sm_count, cudaStream_t s) {
  MegaArgs b = a;
  b.smem_x = (int)decode_mega_smem_floats(a.H, a.QD, a.I, a.q_heads / a.kv_heads, a.head_dim);
  const size_t smem = ((size_t)b.smem_x + MW) * sizeof(float);
  // one block per SM: 512 threads x up to 128 registers fill the register file, so every block of
  // the cooperative grid is resident by construction
  return a.kv_fp32 ? launch_mega_d<float>(b, smem, sm_count, s) : launch_mega_d<uint16_t>(b, smem, sm_count, s);
}

}  // namespace llmc
