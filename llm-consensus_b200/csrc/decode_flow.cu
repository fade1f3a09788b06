// decode_flow.cu -- the batch-1 decode step as one persistent kernel with NO grid barrier.
//
// decode_ring.cu keeps the weight stream running through a per-SM shared-memory ring, but its consumers still meet
// at five grid barriers per layer.  A barrier is a chain of L2 round trips (release the block's stores, arrive,
// poll, acquire) followed by one more to reload the activation vector: 3.7 us of every hand-over in which no SM
// consumes a byte, 25 us per layer against 18.6 us of HBM time.  Here the data IS the flag:
//   * every activation that crosses SMs (q/k/v, the attention output, the residual stream after the O and the down
//     projection, the SwiGLU output, attention split records, per-block argmax candidates) is an 8-byte word
//     (fp32 value, 32-bit tag).  The tag is the running number of the hand-over -- unique per (step, layer, phase),
//     continued across launches -- so a word from any earlier hand-over can never be mistaken for the current one.
//     A reader polls the words it needs; a hand-over costs one store-to-load trip through L2.  8-byte accesses are
//     single-copy atomic, so no fence orders value against tag;
//   * one writer per word: the O / down epilogues add their dot product to the block's own shared-memory copy of
//     the residual stream and store the sum (the same fp32 rounding as the barrier kernel's reduction), so results
//     stay bit-reproducible from run to run;
//   * write-after-read safety needs no extra flag: every phase reads ALL words of its input before it can write one
//     word of its output, and the next writer of a buffer sits at least two such all-to-all hand-overs later;
//   * the K/V row of the new position is written to the page pool by the attention block that owns the last tile
//     (it reads k and v as tagged words), so within a step no block reads pool data another block wrote; visibility
//     of that row to LATER steps is ordered once per step by a fence pair around the argmax hand-over;
//   * producer warp, ring, slot arithmetic, attention tiles and the slot-to-block schedule are decode_ring.cu's.
// Every poll loop gives up after 10 s and traps, so a protocol error surfaces as a CUDA error, not a hung GPU.
// Shapes: as decode_ring.cu (H = 2048, QD and I in {2048, 4096, 8192}).  HBM-bound; algorithmic bytes per step
// as for the other decode paths.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include "attn_tile.cuh"
#include "common.cuh"
#include "kernels.h"
#include "ring_common.cuh"

namespace llmc {

namespace {
using namespace ring;
typedef unsigned long long u64;

constexpr int FLOW_MAX_GRID = 160;   // argmax candidates are polled 5 per lane by one warp

// ---------------------------------------------------------------- tagged words
__device__ __forceinline__ u64 tw_pack(float v, unsigned tag) { return ((u64)tag << 32) | (u64)__float_as_uint(v); }
__device__ __forceinline__ bool tw_ok(u64 w, unsigned tag) { return (unsigned)(w >> 32) == tag; }
__device__ __forceinline__ float tw_val(u64 w) { return __uint_as_float((unsigned)w); }
__device__ __forceinline__ void tw_st1(u64* p, float v, unsigned tag) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(tw_pack(v, tag)) : "memory");
}
__device__ __forceinline__ void tw_st2(u64* p, float v0, float v1, unsigned tag) {   // p 16-byte aligned
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(tw_pack(v0, tag)), "l"(tw_pack(v1, tag))
               : "memory");
}
__device__ __forceinline__ u64 tw_ld1(const u64* p) {
  u64 w;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  return w;
}
__device__ __forceinline__ void tw_ld2(const u64* p, u64& w0, u64& w1) {             // p 16-byte aligned
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
}
// A poll loop calls this once per failed round: after 10 s of waiting the kernel traps instead of hanging.
struct SpinGuard {
  unsigned n = 0;
  u64 t0 = 0;
  __device__ __forceinline__ void miss() {
    if ((++n & 1023u) == 0u) {
      u64 t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      if (t0 == 0) t0 = t;
      else if (t - t0 > 10000000000ull) __trap();
    }
  }
};
__device__ __forceinline__ float tw_wait1(const u64* p, unsigned tag) {
  SpinGuard g;
  u64 w = tw_ld1(p);
  while (!tw_ok(w, tag)) {
    g.miss();
    w = tw_ld1(p);
  }
  return tw_val(w);
}
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

template <int D, int G, typename KVT>
__global__ void __launch_bounds__(RTHREADS, 1) decode_flow_kernel(const MegaArgs a) {
  constexpr int PS = D + 2;
  constexpr int DPL = D / 32;
  using LM = LaneMap<D, G, KVT>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* xraw = reinterpret_cast<float*>(smem_raw + (size_t)NSLOT * SLOT_BYTES);   // the residual stream, this block's copy
  float* att = xraw + SEG;                                                          // attention scratch / activation staging
  __shared__ __align__(8) unsigned long long s_bar[2 * NSLOT];                     // full[NSLOT], empty[NSLOT]
  __shared__ float s_part[2][2][4][2];                                             // [group][buffer][warp][row]
  __shared__ float s_val[RW];
  __shared__ float s_red[RW];
  __shared__ int s_idx[RW];
  __shared__ int s_tok;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int grid = (int)gridDim.x;
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(smem_raw);
  const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(s_bar);
  auto full_bar = [&](int s) { return bar_s + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_s + 8u * (uint32_t)(NSLOT + s); };

  if (tid == 0) {
    for (int s = 0; s < NSLOT; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 4);   // the 4 warps of the group that consumes the slot
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  constexpr int hd2 = D / 2;
  const int H = a.H, QD = a.QD, KD = a.KD, I = a.I, V = a.V;

  // the phase schedule, identical for producer and consumers
  auto phase_desc = [&](int l, int ph) {
    RPhase p;
    p.hd2 = hd2;
    if (l < 0) { p.W = a.lmhead; p.ntasks = V / 2; p.K = H; p.mode = MODE_LMHEAD; return p; }
    const MegaLayer& ly = a.layers[l];
    switch (ph) {
      case 0: p.W = ly.wqkv; p.ntasks = (QD + 2 * KD) / 2; p.K = H; p.mode = MODE_QKV; break;
      case 1: p.W = ly.wo; p.ntasks = H / 2; p.K = QD; p.mode = MODE_RESID; break;
      case 2: p.W = ly.wgu; p.ntasks = I; p.K = H; p.mode = MODE_SWIGLU; break;
      default: p.W = ly.wd; p.ntasks = H / 2; p.K = I; p.mode = MODE_RESID; break;
    }
    return p;
  };

  // ======================================================================== producer (as in decode_ring.cu)
  if (warp >= RW) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    if (warp != RW || lane != 0) return;
    unsigned seq = 0;
    int rot = 0;
    auto produce = [&](const RPhase& p) {
      const int nseg = p.K / SEG, ups = 4 / nseg, nslots = p.ntasks / ups;
      const bool paired = p.mode == MODE_QKV || p.mode == MODE_SWIGLU;
      int first = (int)blockIdx.x - rot;
      if (first < 0) first += grid;
      for (int sl = first; sl < nslots; sl += grid, ++seq) {
        const unsigned r = seq % NSLOT, use = seq / NSLOT;
        if (use > 0) mbar_wait(empty_bar(r), (use - 1) & 1);
        mbar_expect_tx(full_bar(r), SLOT_BYTES);
        const uint32_t dst = ring_s + r * SLOT_BYTES;
        int r0, r1;
        pair_rows(p, sl * ups, r0, r1);
        if (paired) {
          bulk_g2s(dst, p.W + (size_t)r0 * p.K, SLOT_BYTES / 2, full_bar(r));
          bulk_g2s(dst + SLOT_BYTES / 2, p.W + (size_t)r1 * p.K, SLOT_BYTES / 2, full_bar(r));
        } else {
          bulk_g2s(dst, p.W + (size_t)r0 * p.K, SLOT_BYTES, full_bar(r));
        }
      }
      rot = (rot + nslots) % grid;
    };
    for (int step = 0; step < a.nsteps; ++step) {
      for (int l = 0; l < a.L; ++l)
        for (int ph = 0; ph < 4; ++ph) produce(phase_desc(l, ph));
      produce(phase_desc(-1, 0));
    }
    return;
  }

  // ======================================================================== consumers
  asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
  const int group = warp >> 2, wg = warp & 3;
  unsigned seq = 0;
  int rot = 0;
  int tix = 0;
  float xr[SEG / 32];   // this lane's part of the activation segment: k = seg*SEG + c*256 + lane*8 + j
  const bool tracer = a.trace && blockIdx.x == 0 && tid == 0;
  // trace slot tix: [2 tix] = this block has finished the phase before hand-over tix, [2 tix + 1] = it holds the data
  auto trace_arrive = [&]() { if (tracer) a.trace[2 * tix] = gtime(); };
  auto trace_open = [&]() { if (tracer) a.trace[2 * tix + 1] = gtime(); ++tix; };

  // tagged-word buffers (addresses recomputed where used: they would otherwise occupy 14 registers for the whole kernel)
  auto xw_a = [&]() { return a.tw; };                                      // [H]  residual stream after the down projection
  auto xw_b = [&]() { return a.tw + H; };                                  // [H]  residual stream after the O projection
  auto qkvw = [&]() { return a.tw + 2 * H; };                              // [QD + 2 KD]  q (rotated), k (rotated), v
  auto attw = [&]() { return a.tw + 2 * H + QD + 2 * KD; };                // [QD] attention output
  auto actw = [&]() { return a.tw + 2 * H + 2 * QD + 2 * KD; };            // [I]  SwiGLU output
  auto candw = [&]() { return a.tw + 2 * H + 2 * QD + 2 * KD + I; };       // [FLOW_MAX_GRID][2] (value, index) argmax candidates
  auto recw = [&]() { return a.tw + 2 * H + 2 * QD + 2 * KD + I + 2 * FLOW_MAX_GRID; };  // [kv_heads][RING_MAX_SPLIT][G][PS]

  const int slot = a.slot;
  const int pos0 = a.seq_len[slot];
  int tok = a.next_token[slot];
  const int* bt = a.block_table + (size_t)slot * a.bt_stride;
  const float scale = rsqrtf((float)D);
  const int per_step = 5 * a.L + 1;

  // 8 consecutive words per thread (words tid*8 .. tid*8+7 of a 2048-word vector), all four loads in flight
  auto poll8 = [&](const u64* src, unsigned tag, float (&v)[8]) {
    const u64* p = src + tid * 8;
    if (a.flow_poll_mode == 0) {   // one lane per warp watches the warp's first word: 8 pollers per SM, not 256
      if (lane == 0) tw_wait1(p, tag);
      __syncwarp();
    }
    u64 w[8];
    SpinGuard g;
    while (true) {
#pragma unroll
      for (int j = 0; j < 4; ++j) tw_ld2(p + 2 * j, w[2 * j], w[2 * j + 1]);
      bool ok = true;
#pragma unroll
      for (int j = 0; j < 8; ++j) ok = ok && tw_ok(w[j], tag);
      if (ok) break;
      g.miss();
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = tw_val(w[j]);
  };

  // Activation vector -> registers, K == SEG (every warp needs the whole vector): the block polls it ONCE -- 8 words
  // per thread, RMSNorm gain folded, sum of squares reduced through shared memory -- and the warps pick their
  // registers up from shared memory.  is_x: the vector is the residual stream; keep a raw copy for the epilogues.
  // gain != NULL: returns 1/rms.
  auto load_x_full = [&](const u64* src, unsigned tag, const uint16_t* emb_row, const uint16_t* gain, bool is_x) -> float {
    float* xs = att;
    float v[8];
    if (emb_row) {
      const uint4 e = *reinterpret_cast<const uint4*>(emb_row + tid * 8);
      v[0] = bf16lo(e.x); v[1] = bf16hi(e.x); v[2] = bf16lo(e.y); v[3] = bf16hi(e.y);
      v[4] = bf16lo(e.z); v[5] = bf16hi(e.z); v[6] = bf16lo(e.w); v[7] = bf16hi(e.w);
    } else {
      poll8(src, tag, v);
    }
    cbar();   // every warp has finished the previous phase: its staging area and the raw copy may be overwritten
    if (!emb_row) trace_arrive();
    if (is_x) {
      *reinterpret_cast<float4*>(xraw + tid * 8) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(xraw + tid * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (gain) {
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) ss = fmaf(v[j], v[j], ss);
      const uint4 gb = *reinterpret_cast<const uint4*>(gain + tid * 8);
      v[0] *= bf16lo(gb.x); v[1] *= bf16hi(gb.x); v[2] *= bf16lo(gb.y); v[3] *= bf16hi(gb.y);
      v[4] *= bf16lo(gb.z); v[5] *= bf16hi(gb.z); v[6] *= bf16lo(gb.w); v[7] *= bf16hi(gb.w);
      ss = warp_sum(ss);
      if (lane == 0) s_red[warp] = ss;
    }
    // thread (warp, lane) holds k = warp*256 + lane*8 + j: j < 4 -> [warp*256 + lane*4 + j], j >= 4 -> [.. + 128 ..]
    *reinterpret_cast<float4*>(xs + warp * 256 + lane * 4) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(xs + warp * 256 + 128 + lane * 4) = make_float4(v[4], v[5], v[6], v[7]);
    cbar();
#pragma unroll
    for (int c = 0; c < SEG / 256; ++c) {
      const float4 v0 = *reinterpret_cast<const float4*>(xs + c * 256 + lane * 4);
      const float4 v1 = *reinterpret_cast<const float4*>(xs + c * 256 + 128 + lane * 4);
      xr[c * 8 + 0] = v0.x; xr[c * 8 + 1] = v0.y; xr[c * 8 + 2] = v0.z; xr[c * 8 + 3] = v0.w;
      xr[c * 8 + 4] = v1.x; xr[c * 8 + 5] = v1.y; xr[c * 8 + 6] = v1.z; xr[c * 8 + 7] = v1.w;
    }
    float rs = 1.f;
    if (gain) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < RW; ++w) t += s_red[w];
      rs = rsqrtf(t / (float)SEG + a.eps);
    }
    if (!emb_row) trace_open();
    return rs;
  };
  // K > SEG: a warp needs only its own segment (shared with one warp of the other group): its lanes poll it directly,
  // in batches of 8 loads
  auto load_x_seg = [&](const u64* src, unsigned tag, int seg) {
    trace_arrive();
    if (a.flow_poll_mode == 0) {
      if (lane == 0) tw_wait1(src + (size_t)seg * SEG, tag);
      __syncwarp();
    }
#pragma unroll
    for (int c0 = 0; c0 < SEG / 256; c0 += 2) {
      u64 w[16];
      const u64* p = src + (size_t)seg * SEG + c0 * 256 + lane * 8;
      SpinGuard g;
      while (true) {
#pragma unroll
        for (int j = 0; j < 8; ++j) tw_ld2(p + (j >> 2) * 256 + (j & 3) * 2, w[2 * j], w[2 * j + 1]);
        bool ok = true;
#pragma unroll
        for (int j = 0; j < 16; ++j) ok = ok && tw_ok(w[j], tag);
        if (ok) break;
        g.miss();
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) xr[c0 * 8 + j] = tw_val(w[j]);
    }
    trace_open();
  };
  auto load_x_plain = [&](const u64* src, unsigned tag, int K) {
    if (K == SEG) load_x_full(src, tag, nullptr, nullptr, false);
    else load_x_seg(src, tag, wg % (K / SEG));
  };

  // one GEMV phase: this block's slots of the phase, alternating between the two warp groups
  auto run_phase = [&](const RPhase& p, float rs, auto&& epilogue) {
    const int nseg = p.K / SEG, ups = 4 / nseg, nslots = p.ntasks / ups;
    const int uu = wg / nseg, seg = wg - uu * nseg;
    const bool paired = p.mode == MODE_QKV || p.mode == MODE_SWIGLU;
    const uint32_t offA = (uint32_t)((paired ? uu : 2 * uu) * p.K + seg * SEG) * 2u + (uint32_t)lane * 16u;
    const uint32_t offB = (uint32_t)((paired ? ups + uu : 2 * uu + 1) * p.K + seg * SEG) * 2u + (uint32_t)lane * 16u;
    int first = (int)blockIdx.x - rot;
    if (first < 0) first += grid;
    for (int sl = first; sl < nslots; sl += grid, ++seq) {
      if ((int)(seq & 1u) != group) continue;
      const unsigned r = seq % NSLOT, par = (seq / NSLOT) & 1u;
      mbar_wait(full_bar(r), par);
      const uint32_t base = ring_s + r * SLOT_BYTES;
      // two independent chains per row keep the FMA dependency depth at 32
      float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll
      for (int c = 0; c < SEG / 256; ++c) {
        const uint4 wa = lds128(base + offA + c * 512);
        const uint4 wb = lds128(base + offB + c * 512);
        const float* x = xr + c * 8;
        a0 = fmaf(bf16lo(wa.x), x[0], a0); a1 = fmaf(bf16hi(wa.x), x[1], a1);
        a0 = fmaf(bf16lo(wa.y), x[2], a0); a1 = fmaf(bf16hi(wa.y), x[3], a1);
        a0 = fmaf(bf16lo(wa.z), x[4], a0); a1 = fmaf(bf16hi(wa.z), x[5], a1);
        a0 = fmaf(bf16lo(wa.w), x[6], a0); a1 = fmaf(bf16hi(wa.w), x[7], a1);
        b0 = fmaf(bf16lo(wb.x), x[0], b0); b1 = fmaf(bf16hi(wb.x), x[1], b1);
        b0 = fmaf(bf16lo(wb.y), x[2], b0); b1 = fmaf(bf16hi(wb.y), x[3], b1);
        b0 = fmaf(bf16lo(wb.z), x[4], b0); b1 = fmaf(bf16hi(wb.z), x[5], b1);
        b0 = fmaf(bf16lo(wb.w), x[6], b0); b1 = fmaf(bf16hi(wb.w), x[7], b1);
      }
      float acc0 = a0 + a1, acc1 = b0 + b1;
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar(r));   // the slot's bytes are in registers: hand it back
      acc0 = warp_sum(acc0);
      acc1 = warp_sum(acc1);
      if (nseg > 1) {
        const int buf = (seq >> 1) & 1;
        if (lane == 0) {
          s_part[group][buf][wg][0] = acc0;
          s_part[group][buf][wg][1] = acc1;
        }
        gbar(group);
        if (seg == 0) {
          for (int j = 1; j < nseg; ++j) {
            acc0 += s_part[group][buf][wg + j][0];
            acc1 += s_part[group][buf][wg + j][1];
          }
        }
      }
      if (seg == 0) epilogue(sl * ups + uu, acc0 * rs, acc1 * rs);
    }
    rot = (rot + nslots) % grid;
  };

  for (int step = 0; step < a.nsteps; ++step) {
    const int pos = pos0 + step;
    const int ctx = pos + 1;
    const int ntiles = (ctx + TP - 1) / TP;
    const int page_new = bt[pos / a.page_size];
    const int off_new = pos % a.page_size;
    const unsigned tag_step = a.tag0 + (unsigned)step * (unsigned)per_step;
    // this position's RoPE row, spread over the lanes (the QKV epilogue fetches its pair by shuffle)
    constexpr int NR = hd2 > 32 ? hd2 / 32 : 1;
    float rope_c[NR], rope_s[NR];
#pragma unroll
    for (int j = 0; j < NR; ++j) {
      const int i = (j * 32 + lane) % hd2;
      rope_c[j] = a.rope_cos[(size_t)pos * hd2 + i];
      rope_s[j] = a.rope_sin[(size_t)pos * hd2 + i];
    }

    for (int l = 0; l < a.L; ++l) {
      const MegaLayer& ly = a.layers[l];
      KVT* kpool = reinterpret_cast<KVT*>(ly.kp);
      KVT* vpool = reinterpret_cast<KVT*>(ly.vp);
      const unsigned t_qkv = tag_step + 5u * (unsigned)l, t_att = t_qkv + 1, t_o = t_qkv + 2, t_gu = t_qkv + 3,
                     t_d = t_qkv + 4;
      // splits per kv head: about attn_tiles_per_block tiles per block, bounded by the record layout and the grid
      const int nsp = max(1, min(min(RING_MAX_SPLIT, grid / a.kv_heads),
                                 (ntiles + a.attn_tiles_per_block - 1) / a.attn_tiles_per_block));

      // ---------------- P1: RMSNorm -> QKV -> RoPE, results as tagged words
      {
        const float rs = load_x_full(xw_a(), t_qkv - 1, l == 0 ? a.emb + (size_t)tok * H : nullptr, ly.g1, true);
        run_phase(phase_desc(l, 0), rs, [&](int g, float v0, float v1) {
          const int h = g / hd2, i = g - h * hd2;
          float cv = rope_c[0], sv = rope_s[0];
          if constexpr (NR > 1) {
            if (i >= 32) { cv = rope_c[1]; sv = rope_s[1]; }
          }
          const float c = __shfl_sync(0xffffffffu, cv, i & 31);
          const float s = __shfl_sync(0xffffffffu, sv, i & 31);
          if (lane == 0) {
            if (h < a.q_heads + a.kv_heads) {
              const float o0 = v0 * c - v1 * s, o1 = v1 * c + v0 * s;
              v0 = o0; v1 = o1;
            }
            // fused head order: q heads, k heads, v heads
            tw_st1(qkvw() + (size_t)h * D + i, v0, t_qkv);
            tw_st1(qkvw() + (size_t)h * D + i + hd2, v1, t_qkv);
          }
        });
        // The K/V tiles this block's warps will read in the attention phase were written by earlier steps and have
        // long been evicted from L2 by the weight stream: ask L2 for them now, so that the tile loads of the
        // attention phase cost an L2 trip instead of an HBM trip.
        if ((int)blockIdx.x < a.kv_heads * nsp) {
          const int kvh = blockIdx.x % a.kv_heads, split = blockIdx.x / a.kv_heads;
          constexpr int TILE_LINES = TP * D * (int)sizeof(KVT) / 128;
          for (int tile = split + nsp * warp; tile < ntiles; tile += nsp * RW) {
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

      // ---------------- P2: attention: block b < kv_heads * nsp takes (kv head, split); the others go straight on
      if ((int)blockIdx.x < a.kv_heads * nsp) {
        float* qs = att;                                           // [G][D]
        float* part = att + G * D;                                 // [RW][G][PS]
        float* pw = part + (size_t)RW * G * PS + (size_t)warp * G * (TP + 1);
        const int kvh = blockIdx.x % a.kv_heads, split = blockIdx.x / a.kv_heads;
        const bool owner = ((ntiles - 1) % nsp) == split;          // this block's tiles include the new position
        // q of the GQA group, and -- in the block that owns the last tile -- the new k and v rows
        constexpr int QPT = (G * D + RW * 32 - 1) / (RW * 32);
        float qv[QPT];
#pragma unroll
        for (int j = 0; j < QPT; ++j) {
          const int i = tid + j * RW * 32;
          qv[j] = (i < G * D) ? tw_wait1(qkvw() + (size_t)kvh * G * D + i, t_qkv) : 0.f;
        }
        float kvn = 0.f;
        if (owner && tid < 2 * D)
          kvn = tw_wait1(qkvw() + (tid < D ? QD + kvh * D + tid : QD + KD + kvh * D + (tid - D)), t_qkv);
        cbar();
        trace_arrive();
#pragma unroll
        for (int j = 0; j < QPT; ++j) {
          const int i = tid + j * RW * 32;
          if (i < G * D) qs[i] = qv[j];
        }
        if (owner && tid < 2 * D) {
          KVT* dst = (tid < D ? kpool : vpool) + (((size_t)page_new * a.kv_heads + kvh) * a.page_size + off_new) * D +
                     (tid < D ? tid : tid - D);
          if constexpr (sizeof(KVT) == 4) *dst = kvn;
          else *dst = f32_to_bf16_bits(kvn);
        }
        cbar();   // q staged; the new K/V row is visible to this block's tile loads
        trace_open();
        const int c = lane % LM::LPR;
        float m[G], lsum[G], o[G][LM::EPL];
#pragma unroll
        for (int h = 0; h < G; ++h) {
          m[h] = -INFINITY;
          lsum[h] = 0.f;
#pragma unroll
          for (int e = 0; e < LM::EPL; ++e) o[h][e] = 0.f;
        }
        for (int tile = split + nsp * warp; tile < ntiles; tile += nsp * RW) {
          const int p0 = tile * TP;
          const int valid = min(TP, ctx - p0);
          const int page = bt[p0 / a.page_size];
          const size_t base = (((size_t)page * a.kv_heads + kvh) * a.page_size + (p0 % a.page_size)) * D;
          attn_tile_step<D, G, KVT, true, 16>(qs, pw, kpool + base + c * LM::EPV, vpool + base + c * LM::EPV, valid,
                                              scale, m, lsum, o);
        }
        attn_write_record<D, G, KVT>(part + (size_t)warp * G * PS, m, lsum, o);
        cbar();
        if (warp < G) {  // warp h merges head h over the block's warps
          const int h = warp;
          float M = -INFINITY;
#pragma unroll
          for (int w = 0; w < RW; ++w) M = fmaxf(M, part[((size_t)w * G + h) * PS + D]);
          float L = 0.f, O[DPL];
#pragma unroll
          for (int i = 0; i < DPL; ++i) O[i] = 0.f;
          if (M != -INFINITY) {
#pragma unroll
            for (int w = 0; w < RW; ++w) {
              const float* pp = part + ((size_t)w * G + h) * PS;
              const float mw = pp[D];
              const float wt = (mw == -INFINITY) ? 0.f : __expf(mw - M);
              L = fmaf(pp[D + 1], wt, L);
#pragma unroll
              for (int i = 0; i < DPL; ++i) O[i] = fmaf(pp[lane * DPL + i], wt, O[i]);
            }
          }
          if (split > 0) {
            // this split's record for head h: D output words (unnormalised), then (m, l)
            u64* rec = recw() + (((size_t)kvh * RING_MAX_SPLIT + split) * G + h) * PS;
            if constexpr (DPL % 2 == 0) {
#pragma unroll
              for (int i = 0; i < DPL; i += 2) tw_st2(rec + lane * DPL + i, O[i], O[i + 1], t_att);
            } else {
#pragma unroll
              for (int i = 0; i < DPL; ++i) tw_st1(rec + lane * DPL + i, O[i], t_att);
            }
            if (lane == 0) tw_st2(rec + D, M, L, t_att);
          } else {
            // split 0 folds the other splits' records into its own state as they arrive, 8 at a time
            for (int s0 = 1; s0 < nsp; s0 += 8) {
              u64 wml[8][2], wo[8][DPL];
              unsigned live = 0u;
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (s0 + j < nsp) live |= 1u << j;
              SpinGuard g;
              while (true) {
                bool ok = true;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (live & (1u << j)) {
                    const u64* rec = recw() + (((size_t)kvh * RING_MAX_SPLIT + s0 + j) * G + h) * PS;
                    tw_ld2(rec + D, wml[j][0], wml[j][1]);
                    if constexpr (DPL % 2 == 0) {
#pragma unroll
                      for (int i = 0; i < DPL; i += 2) tw_ld2(rec + lane * DPL + i, wo[j][i], wo[j][i + 1]);
                    } else {
#pragma unroll
                      for (int i = 0; i < DPL; ++i) wo[j][i] = tw_ld1(rec + lane * DPL + i);
                    }
                  }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (live & (1u << j)) {
                    ok = ok && tw_ok(wml[j][0], t_att) && tw_ok(wml[j][1], t_att);
#pragma unroll
                    for (int i = 0; i < DPL; ++i) ok = ok && tw_ok(wo[j][i], t_att);
                  }
                if (__all_sync(0xffffffffu, ok)) break;
                g.miss();
              }
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (live & (1u << j)) {
                  const float ms = tw_val(wml[j][0]), ls = tw_val(wml[j][1]);
                  const float Mn = fmaxf(M, ms);
                  const float w0 = (M == -INFINITY) ? 0.f : __expf(M - Mn);
                  const float w1 = (ms == -INFINITY) ? 0.f : __expf(ms - Mn);
                  L = L * w0 + ls * w1;
#pragma unroll
                  for (int i = 0; i < DPL; ++i) O[i] = O[i] * w0 + tw_val(wo[j][i]) * w1;
                  M = Mn;
                }
            }
          }
          if (split == 0) {
            const float inv = 1.0f / L;
            u64* dst = attw() + (size_t)(kvh * G + h) * D + lane * DPL;
            if constexpr (DPL % 2 == 0) {
#pragma unroll
              for (int i = 0; i < DPL; i += 2) tw_st2(dst + i, O[i] * inv, O[i + 1] * inv, t_att);
            } else {
#pragma unroll
              for (int i = 0; i < DPL; ++i) tw_st1(dst + i, O[i] * inv, t_att);
            }
          }
        }
      } else {
        // keep the trace slots aligned for blocks without attention work (block 0 always has some)
        ++tix;
      }

      // ---------------- P3: x += Wo attn
      {
        const RPhase p = phase_desc(l, 1);
        load_x_plain(attw(), t_att, p.K);
        run_phase(p, 1.f, [&](int task, float v0, float v1) {
          if (lane == 0) tw_st2(xw_b() + 2 * task, xraw[2 * task] + v0, xraw[2 * task + 1] + v1, t_o);
        });
      }

      // ---------------- P4: act = silu(Wg h) * (Wu h), h = RMSNorm(x)
      {
        const float rs = load_x_full(xw_b(), t_o, nullptr, ly.g2, true);
        run_phase(phase_desc(l, 2), rs, [&](int g, float gt, float up) {
          if (lane == 0) tw_st1(actw() + g, gt / (1.0f + __expf(-gt)) * up, t_gu);
        });
      }

      // ---------------- P5: x += Wd act
      {
        const RPhase p = phase_desc(l, 3);
        load_x_plain(actw(), t_gu, p.K);
        run_phase(p, 1.f, [&](int task, float v0, float v1) {
          if (lane == 0) tw_st2(xw_a() + 2 * task, xraw[2 * task] + v0, xraw[2 * task + 1] + v1, t_d);
        });
      }
    }

    // ---------------- LM head + greedy argmax (lowest index wins ties)
    {
      const unsigned t_lm = tag_step + 5u * (unsigned)a.L;
      const float rs = load_x_full(xw_a(), t_lm - 1, nullptr, a.gf, true);
      float best = -INFINITY;
      int bi = 0x7fffffff;
      const bool store_logits = a.logits != nullptr && step == a.nsteps - 1;
      run_phase(phase_desc(-1, 0), rs, [&](int g, float v0, float v1) {
        if (store_logits && lane == 0) *reinterpret_cast<float2*>(a.logits + 2 * g) = make_float2(v0, v1);
        if (v0 > best || (v0 == best && 2 * g < bi)) { best = v0; bi = 2 * g; }
        if (v1 > best || (v1 == best && 2 * g + 1 < bi)) { best = v1; bi = 2 * g + 1; }
      });
      if (lane == 0) {
        s_val[warp] = best;
        s_idx[warp] = bi;
      }
      cbar();
      if (tid == 0) {
#pragma unroll
        for (int w = 1; w < RW; ++w)
          if (s_val[w] > best || (s_val[w] == best && s_idx[w] < bi)) { best = s_val[w]; bi = s_idx[w]; }
        // once per step: everything this block wrote so far (the K/V rows of this step among it) is ordered before
        // the candidate, and each reader fences after it has seen all candidates
        fence_gpu();
        tw_st2(candw() + 2 * blockIdx.x, best, __int_as_float(bi), t_lm);
      }
      // every block reduces the per-block candidates itself
      if (warp == 0) {
        trace_arrive();
        constexpr int NCL = FLOW_MAX_GRID / 32;
        u64 wv[NCL], wi[NCL];
        unsigned live = 0u;
#pragma unroll
        for (int j = 0; j < NCL; ++j)
          if (lane + 32 * j < grid) live |= 1u << j;
        SpinGuard g;
        while (true) {
          bool ok = true;
#pragma unroll
          for (int j = 0; j < NCL; ++j)
            if (live & (1u << j)) {
              tw_ld2(candw() + 2 * (lane + 32 * j), wv[j], wi[j]);
              ok = ok && tw_ok(wv[j], t_lm) && tw_ok(wi[j], t_lm);
            }
          if (ok) break;
          g.miss();
        }
        float bv = -INFINITY;
        int bx = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < NCL; ++j)
          if (live & (1u << j)) {
            const float v = tw_val(wv[j]);
            const int ix = __float_as_int(tw_val(wi[j]));
            if (v > bv || (v == bv && ix < bx)) { bv = v; bx = ix; }
          }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bx, o);
          if (ov > bv || (ov == bv && oi < bx)) { bv = ov; bx = oi; }
        }
        fence_gpu();
        if (lane == 0) {
          s_tok = bx;
          if (blockIdx.x == 0) a.out_tokens[step] = bx;
        }
        trace_open();
      } else {
        ++tix;
      }
      cbar();
      tok = s_tok;
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    a.next_token[slot] = tok;
    a.seq_len[slot] = pos0 + a.nsteps;
  }
}

size_t flow_smem_bytes(int G, int D) {
  size_t att = (size_t)G * D + (size_t)RW * G * (D + 2) + (size_t)RW * G * (TP + 1);
  if (att < (size_t)SEG) att = SEG;   // the same region stages the activation vector outside the attention phase
  return (size_t)NSLOT * SLOT_BYTES + (size_t)SEG * sizeof(float) + att * sizeof(float);
}

template <int D, int G, typename KVT>
cudaError_t launch_flow_t(const MegaArgs& a, int grid, cudaStream_t s) {
  auto kern = decode_flow_kernel<D, G, KVT>;
  const size_t smem = flow_smem_bytes(G, D);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  MegaArgs copy = a;
  void* params[] = {&copy};
  // cooperative launch: every block must be resident, the blocks wait for each other's words
  return cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3(RTHREADS), params, smem, s);
}

template <int D, typename KVT>
cudaError_t launch_flow_g(int G, const MegaArgs& a, int grid, cudaStream_t s) {
  switch (G) {
    case 1: return launch_flow_t<D, 1, KVT>(a, grid, s);
    case 2: return launch_flow_t<D, 2, KVT>(a, grid, s);
    case 4: return launch_flow_t<D, 4, KVT>(a, grid, s);
    case 8: return launch_flow_t<D, 8, KVT>(a, grid, s);
  }
  return cudaErrorInvalidValue;
}

template <typename KVT>
cudaError_t launch_flow_d(const MegaArgs& a, int grid, cudaStream_t s) {
  const int G = a.q_heads / a.kv_heads;
  switch (a.head_dim) {
    case 32: return launch_flow_g<32, KVT>(G, a, grid, s);
    case 64: return launch_flow_g<64, KVT>(G, a, grid, s);
    case 128: return launch_flow_g<128, KVT>(G, a, grid, s);
  }
  return cudaErrorInvalidValue;
}

bool seg_ok(int K) { return K == SEG || K == 2 * SEG || K == 4 * SEG; }

}  // namespace

bool decode_flow_supported(int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim, int sm_count) {
  if (kv_heads < 1 || q_heads % kv_heads) return false;
  const int G = q_heads / kv_heads;
  if (!(G == 1 || G == 2 || G == 4 || G == 8)) return false;
  if (!(head_dim == 32 || head_dim == 64 || head_dim == 128)) return false;
  if (H != ring::SEG || !seg_ok(QD) || !seg_ok(I)) return false;
  if (((QD + 2 * KD) / 2) % 4 || V % 8 || I % 64) return false;
  if (sm_count < kv_heads || sm_count > FLOW_MAX_GRID) return false;
  return flow_smem_bytes(G, head_dim) + 1024 <= 227 * 1024;
}

size_t decode_flow_words(int H, int QD, int KD, int I, int q_heads, int kv_heads, int head_dim, int sm_count) {
  (void)sm_count;
  const int G = q_heads / kv_heads;
  return (size_t)2 * H + (size_t)(QD + 2 * KD) + (size_t)QD + (size_t)I + (size_t)2 * FLOW_MAX_GRID +
         (size_t)kv_heads * ring::RING_MAX_SPLIT * G * (head_dim + 2);
}

cudaError_t launch_decode_flow(const MegaArgs& a, int sm_count, cudaStream_t s) {
  if (!decode_flow_supported(a.H, a.QD, a.KD, a.I, a.V, a.q_heads, a.kv_heads, a.head_dim, sm_count) || !a.tw || !a.tag0)
    return cudaErrorInvalidValue;
  return a.kv_fp32 ? launch_flow_d<float>(a, sm_count, s) : launch_flow_d<uint16_t>(a, sm_count, s);
}

}  // namespace llmc
