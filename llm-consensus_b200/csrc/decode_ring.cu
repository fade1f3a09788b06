// decode_ring.cu -- the batch-1 decode step as one persistent kernel whose weight stream never stops.
//
// decode_mega.cu already runs all phases of all steps inside one cooperative kernel, but every warp still
// pulls its own weights through registers, so the stream restarts after each of the 81 grid barriers and the
// small phases (QKV, O) run at a quarter of the HBM rate (profiles/README.md, per-phase trace).  Here the
// weights travel through a per-SM shared-memory ring instead:
//   * one producer thread per block walks the whole schedule -- every phase of every layer of every step --
//     and issues `cp.async.bulk` (TMA, 1-D) copies of 32 KB slots into a 6-slot ring, throttled only by the
//     ring's empty barriers.  Weights never depend on activations, so the producer runs through grid barriers
//     and through the attention phase: 148 x 192 KB = 28 MB (4 us of HBM time) is always in flight or on chip;
//   * 8 consumer warps (two groups of 4 that alternate slots) wait on a slot's full barrier, read the rows with
//     conflict-free 16-byte shared loads and multiply with the activation vector they hold in REGISTERS (a warp
//     owns a 2048-element k segment = 64 registers per lane).  The block reads the vector from L2 once per phase
//     and redistributes it through shared memory, RMSNorm statistics included;
//   * the block is three full warpgroups (8 consumer warps, 4 producer-side warps of which one thread works), so
//     that setmaxnreg can hand the producer side's registers to the consumers (24 / 232 per thread);
//   * a slot is 8 row segments: 4 row pairs (RoPE pairs / gate-up pairs / adjacent rows) when K = 2048, 2 when
//     K = 4096, 1 pair split over 4 warps when K = 8192.  Slots are dealt round-robin to the blocks with a
//     rotation that continues across phases, so the per-block byte count is balanced over a layer;
//   * activations cross SMs through L2 and phases are separated by the same grid barrier as in decode_mega.cu;
//     the attention phase is that kernel's (block = (kv head, split), warps = 32-position tiles), with up to 16
//     splits per kv head and the step's K/V tiles requested into L2 before the QKV barrier.
// Shapes: H = 2048, QD and I in {2048, 4096, 8192} (llama-3.2-1b); other shapes keep decode_mega.cu.
// HBM-bound; algorithmic bytes per step as for the other decode paths.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include <type_traits>

#include "attn_tile.cuh"
#include "common.cuh"
#include "kernels.h"
#include "ring_common.cuh"

namespace llmc {

namespace {
using namespace ring;

template <int D, int G, typename KVT>
__global__ void __launch_bounds__(RTHREADS, 1) decode_ring_kernel(const MegaArgs a) {
  constexpr int PS = D + 2;
  using LM = LaneMap<D, G, KVT>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* att = reinterpret_cast<float*>(smem_raw + (size_t)NSLOT * SLOT_BYTES);   // attention phase scratch
  __shared__ __align__(8) unsigned long long s_bar[2 * NSLOT];                     // full[NSLOT], empty[NSLOT]
  __shared__ float s_part[2][2][4][2];                                             // [group][buffer][warp][row]
  __shared__ float s_val[RW];
  __shared__ float s_red[RW];
  __shared__ int s_idx[RW];
  __shared__ int s_tok;
  __shared__ int s_last;

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

  // ======================================================================== producer
  // Register budget: 9 warps would put three on one scheduler and cap every thread at 168 registers.  With a full
  // third warpgroup the producer side hands its registers back (setmaxnreg needs whole warpgroups) and the eight
  // consumer warps get 232 each, which the attention phase uses to request a whole K / V tile at once.
  if (warp >= RW) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    if (warp != RW || lane != 0) return;
    unsigned seq = 0;
    int rot = 0;
    const bool stats = a.trace && blockIdx.x == 0;
    long long t_wait = 0, t_begin = clock64();
    auto produce = [&](const RPhase& p) {
      const int nseg = p.K / SEG, ups = 4 / nseg, nslots = p.ntasks / ups;
      const bool paired = p.mode == MODE_QKV || p.mode == MODE_SWIGLU;
      int first = (int)blockIdx.x - rot;
      if (first < 0) first += grid;
      for (int sl = first; sl < nslots; sl += grid, ++seq) {
        const unsigned r = seq % NSLOT, use = seq / NSLOT;
        if (use > 0) {
          const long long t0 = stats ? clock64() : 0;
          mbar_wait(empty_bar(r), (use - 1) & 1);
          if (stats) t_wait += clock64() - t0;
        }
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
    if (stats) {
      unsigned long long* st = a.trace + (size_t)2 * (5 * a.L + 1) * a.nsteps;
      st[8] = (unsigned long long)t_wait;
      st[9] = (unsigned long long)(clock64() - t_begin);
      st[10] = seq;
    }
    return;
  }

  // ======================================================================== consumers
  asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
  const int group = warp >> 2, wg = warp & 3;
  unsigned seq = 0;
  int rot = 0;
  unsigned bar_target = gridDim.x;
  int tix = 0;
  float xr[SEG / 32];   // this lane's part of the activation segment: k = seg*SEG + c*256 + lane*8 + j
  const bool stats = a.trace && blockIdx.x == 0 && warp == 0;
  long long t_full = 0, t_x = 0, t_bar = 0, t_slots = 0;
  const long long t_begin = clock64();

  const int slot = a.slot;
  const int pos0 = a.seq_len[slot];
  int tok = a.next_token[slot];
  const int* bt = a.block_table + (size_t)slot * a.bt_stride;
  const float scale = rsqrtf((float)D);

  auto grid_barrier = [&]() {
    const long long tb0 = stats ? clock64() : 0;
    cbar();
    if (tid == 0) {
      if (a.trace && blockIdx.x == 0) a.trace[2 * tix] = gtime();
      // LLMC_MEGA_TRACE: every block's arrival time at the barriers of the second step's layer 1 (5 barriers), to
      // see how far apart the blocks finish a phase
      if (a.trace && a.nsteps > 1 && tix >= (5 * a.L + 1) + 5 && tix < (5 * a.L + 1) + 10)
        a.trace[(size_t)2 * (5 * a.L + 1) * a.nsteps + 16 + (size_t)(tix - (5 * a.L + 1) - 5) * gridDim.x + blockIdx.x] = gtime();
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.barrier) : "memory");
      unsigned v;
      if (a.ring_poll_mode == 1) {          // relaxed spin, one acquire load once the count is there (default: the
                                            // acquire-load spin costs 0.5 % of the step; LLMC_RING_POLL=0 selects it)
        do {
          asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.barrier) : "memory");
        } while (v < bar_target);
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.barrier) : "memory");
      } else {
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.barrier) : "memory");
        } while (v < bar_target);
      }
      if (a.trace && blockIdx.x == 0) a.trace[2 * tix + 1] = gtime();
    }
    cbar();
    if (stats) t_bar += clock64() - tb0;
    bar_target += gridDim.x;
    ++tix;
  };

  // Activation vector -> registers.
  // K == SEG (every warp needs the whole vector): the block reads it from L2 ONCE -- 8 floats per thread, RMSNorm
  // gain folded, sum of squares reduced through shared memory -- and the warps pick their registers up from
  // shared memory.  (Each warp loading the vector itself costs 2.6 us per phase: 148 x 8 warps hammer the same
  // 64 L2 lines right after the barrier.)  gain != NULL: returns 1/rms.
  auto load_x_full = [&](const float* src, const uint16_t* emb_row, const uint16_t* gain, float* x_out) -> float {
    const long long tx0 = stats ? clock64() : 0;
    float* xs = att;   // the attention scratch is free outside the attention phase
    {
      const int k = tid * 8;
      float4 v0, v1;
      if (emb_row) {
        const uint4 e = *reinterpret_cast<const uint4*>(emb_row + k);
        v0 = make_float4(bf16lo(e.x), bf16hi(e.x), bf16lo(e.y), bf16hi(e.y));
        v1 = make_float4(bf16lo(e.z), bf16hi(e.z), bf16lo(e.w), bf16hi(e.w));
        if (x_out) {
          *reinterpret_cast<float4*>(x_out + k) = v0;
          *reinterpret_cast<float4*>(x_out + k + 4) = v1;
        }
      } else {
        v0 = __ldcg(reinterpret_cast<const float4*>(src + k));
        v1 = __ldcg(reinterpret_cast<const float4*>(src + k + 4));
      }
      if (gain) {
        float ss = v0.x * v0.x + v0.y * v0.y + v0.z * v0.z + v0.w * v0.w;
        ss += v1.x * v1.x + v1.y * v1.y + v1.z * v1.z + v1.w * v1.w;
        const uint4 gb = *reinterpret_cast<const uint4*>(gain + k);
        v0.x *= bf16lo(gb.x); v0.y *= bf16hi(gb.x); v0.z *= bf16lo(gb.y); v0.w *= bf16hi(gb.y);
        v1.x *= bf16lo(gb.z); v1.y *= bf16hi(gb.z); v1.z *= bf16lo(gb.w); v1.w *= bf16hi(gb.w);
        ss = warp_sum(ss);
        if (lane == 0) s_red[warp] = ss;
      }
      // thread (warp, lane) holds k = warp*256 + lane*8 + j: j < 4 -> [warp*256 + lane*4 + j], j >= 4 -> [.. + 128 ..]
      *reinterpret_cast<float4*>(xs + warp * 256 + lane * 4) = v0;
      *reinterpret_cast<float4*>(xs + warp * 256 + 128 + lane * 4) = v1;
    }
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
    if (stats) t_x += clock64() - tx0 + (long long)(xr[0] == 123.456f);   // the compare keeps the loads inside the window
    return rs;
  };
  // K > SEG: a warp needs only its own segment (shared with one warp of the other group): straight from L2
  auto load_x_seg = [&](const float* src, int seg) {
    const long long tx0 = stats ? clock64() : 0;
#pragma unroll
    for (int c = 0; c < SEG / 256; ++c) {
      const int k = seg * SEG + c * 256 + lane * 8;
      const float4 v0 = __ldcg(reinterpret_cast<const float4*>(src + k));
      const float4 v1 = __ldcg(reinterpret_cast<const float4*>(src + k + 4));
      xr[c * 8 + 0] = v0.x; xr[c * 8 + 1] = v0.y; xr[c * 8 + 2] = v0.z; xr[c * 8 + 3] = v0.w;
      xr[c * 8 + 4] = v1.x; xr[c * 8 + 5] = v1.y; xr[c * 8 + 6] = v1.z; xr[c * 8 + 7] = v1.w;
    }
    if (stats) t_x += clock64() - tx0 + (long long)(xr[0] == 123.456f);
  };
  auto load_x_plain = [&](const float* src, int K) {
    if (K == SEG) load_x_full(src, nullptr, nullptr, nullptr);
    else load_x_seg(src, wg % (K / SEG));
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
      const long long tf0 = stats ? clock64() : 0;
      mbar_wait(full_bar(r), par);
      const long long tf1 = stats ? clock64() : 0;
      if (stats) t_full += tf1 - tf0;
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
      if (stats) t_slots += clock64() - tf1;
    }
    rot = (rot + nslots) % grid;
  };
  auto resid_epilogue = [&](int task, float v0, float v1) {
    // one writer per row: a fire-and-forget reduction (RED) gives the same fp32 sum as load-add-store
    // without a round trip to L2 in the slot loop
    if (lane == 0) {
      atomicAdd(a.x + 2 * task, v0);
      atomicAdd(a.x + 2 * task + 1, v1);
    }
  };

  for (int step = 0; step < a.nsteps; ++step) {
    const int pos = pos0 + step;
    const int ctx = pos + 1;
    const int ntiles = (ctx + TP - 1) / TP;
    const int page_new = bt[pos / a.page_size];
    const int off_new = pos % a.page_size;
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
      const int rec_stride = min(RING_MAX_SPLIT, a.max_tiles);   // split records per kv head in the scratch
      // splits per kv head: about attn_tiles_per_block tiles per block, bounded by the scratch layout, the scratch
      // size (max_tiles records per q head) and the number of blocks
      const int nsp = max(1, min(min(RING_MAX_SPLIT, min(a.max_tiles, grid / a.kv_heads)),
                                 (ntiles + a.attn_tiles_per_block - 1) / a.attn_tiles_per_block));

      // ---------------- P1: RMSNorm -> QKV -> RoPE -> KV append
      {
        const float rs = load_x_full(a.x, l == 0 ? a.emb + (size_t)tok * H : nullptr, ly.g1,
                                     (l == 0 && blockIdx.x == 0) ? a.x : nullptr);
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
        // The K/V tiles this block's warps will read in the attention phase were written by earlier steps (only the
        // newest row is from this one) and have long been evicted from L2 by the weight stream: ask L2 for them now,
        // so the two dependent tile loads (K, then V) of the attention phase cost an L2 trip instead of an HBM trip.
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
        grid_barrier();
      }

      // ---------------- P2: attention (as in decode_mega.cu): block b < kv_heads * nsp takes (kv head, split)
      {
        float* qs = att;                                           // [G][D]
        float* part = att + G * D;                                 // [RW][G][PS]
        float* pw = part + (size_t)RW * G * PS + (size_t)warp * G * (TP + 1);
        if ((int)blockIdx.x < a.kv_heads * nsp) {
          const int kvh = blockIdx.x % a.kv_heads, split = blockIdx.x / a.kv_heads;
          for (int i = tid; i < G * D; i += RW * 32) qs[i] = __ldcg(a.q + (size_t)kvh * G * D + i);
          cbar();
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
            constexpr int DPL = D / 32;
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
            if (nsp == 1) {   // the block saw the whole context: done, no cross-block merge
              const float inv = 1.0f / L;
#pragma unroll
              for (int i = 0; i < DPL; ++i) a.attn[(size_t)(kvh * G + h) * D + lane * DPL + i] = O[i] * inv;
            } else {
              float* rec = a.attn_scratch + (((size_t)kvh * rec_stride + split) * G + h) * PS;
#pragma unroll
              for (int i = 0; i < DPL; ++i) rec[lane * DPL + i] = O[i];
              if (lane == 0) {
                rec[D] = M;
                rec[D + 1] = L;
              }
            }
          }
          // the last block to finish this kv head merges its nsp split records
          if (nsp > 1) {
            __threadfence();
            cbar();
            if (tid == 0) s_last = (atomicAdd(a.attn_counters + kvh, 1u) == (unsigned)(nsp - 1)) ? 1 : 0;
            cbar();
          }
          if (nsp > 1 && s_last && warp < G) {
            __threadfence();
            constexpr int DPL = D / 32;
            const int h = warp;
            const float* rb = a.attn_scratch + ((size_t)kvh * rec_stride * G + h) * PS;
            // all record loads of a lane are issued together (one L2 round trip); the common case of at most 8
            // splits runs the half-size copy of the code
            auto merge = [&](auto cap) {
              constexpr int CAP = decltype(cap)::value;
              float mv[CAP], lv[CAP], ov[CAP][DPL];
#pragma unroll
              for (int sp = 0; sp < CAP; ++sp) {
                const float* r = rb + (size_t)sp * G * PS;
                const bool live = sp < nsp;
                mv[sp] = live ? __ldcg(r + D) : -INFINITY;
                lv[sp] = live ? __ldcg(r + D + 1) : 0.f;
#pragma unroll
                for (int i = 0; i < DPL; ++i) ov[sp][i] = live ? __ldcg(r + lane * DPL + i) : 0.f;
              }
              float M = -INFINITY;
#pragma unroll
              for (int sp = 0; sp < CAP; ++sp) M = fmaxf(M, mv[sp]);
              float L = 0.f, O[DPL];
#pragma unroll
              for (int i = 0; i < DPL; ++i) O[i] = 0.f;
#pragma unroll
              for (int sp = 0; sp < CAP; ++sp) {
                const float wt = (mv[sp] == -INFINITY) ? 0.f : __expf(mv[sp] - M);
                L = fmaf(lv[sp], wt, L);
#pragma unroll
                for (int i = 0; i < DPL; ++i) O[i] = fmaf(ov[sp][i], wt, O[i]);
              }
              const float inv = 1.0f / L;
#pragma unroll
              for (int i = 0; i < DPL; ++i) a.attn[(size_t)(kvh * G + h) * D + lane * DPL + i] = O[i] * inv;
            };
            if (nsp <= 8) merge(std::integral_constant<int, 8>{});
            else merge(std::integral_constant<int, RING_MAX_SPLIT>{});
            if (lane == 0 && h == 0) a.attn_counters[kvh] = 0u;
          }
        }
        grid_barrier();
      }

      // ---------------- P3: x += Wo attn
      {
        const RPhase p = phase_desc(l, 1);
        load_x_plain(a.attn, p.K);
        run_phase(p, 1.f, resid_epilogue);
        grid_barrier();
      }

      // ---------------- P4: act = silu(Wg h) * (Wu h), h = RMSNorm(x)
      {
        const float rs = load_x_full(a.x, nullptr, ly.g2, nullptr);
        run_phase(phase_desc(l, 2), rs, [&](int g, float gt, float up) {
          if (lane == 0) a.act[g] = gt / (1.0f + __expf(-gt)) * up;
        });
        grid_barrier();
      }

      // ---------------- P5: x += Wd act
      {
        const RPhase p = phase_desc(l, 3);
        load_x_plain(a.act, p.K);
        run_phase(p, 1.f, resid_epilogue);
        grid_barrier();
      }
    }

    // ---------------- LM head + greedy argmax (lowest index wins ties)
    {
      const float rs = load_x_full(a.x, nullptr, a.gf, nullptr);
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
        a.part_val[blockIdx.x] = best;
        a.part_idx[blockIdx.x] = bi;
      }
      grid_barrier();
      // every block reduces the per-block candidates itself: no second exchange
      if (warp == 0) {
        float bv = -INFINITY;
        int bx = 0x7fffffff;
        for (int j = lane; j < grid; j += 32) {
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
      cbar();
      tok = s_tok;
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    a.next_token[slot] = tok;
    a.seq_len[slot] = pos0 + a.nsteps;
  }
  if (stats && lane == 0) {   // debug cycle counters of block 0 / warp 0 (LLMC_MEGA_TRACE=1)
    unsigned long long* st = a.trace + (size_t)2 * (5 * a.L + 1) * a.nsteps;
    st[0] = (unsigned long long)t_full;
    st[1] = (unsigned long long)t_x;
    st[2] = (unsigned long long)t_slots;
    st[3] = (unsigned long long)t_bar;
    st[4] = (unsigned long long)(clock64() - t_begin);
    st[5] = seq;
  }
}

size_t ring_smem_bytes(int G, int D) {
  size_t att = (size_t)G * D + (size_t)RW * G * (D + 2) + (size_t)RW * G * (TP + 1);
  if (att < (size_t)SEG) att = SEG;   // the same region stages the activation vector outside the attention phase
  return (size_t)NSLOT * SLOT_BYTES + att * sizeof(float);
}

template <int D, int G, typename KVT>
cudaError_t launch_ring_t(const MegaArgs& a, int grid, cudaStream_t s) {
  auto kern = decode_ring_kernel<D, G, KVT>;
  const size_t smem = ring_smem_bytes(G, D);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  MegaArgs copy = a;
  void* params[] = {&copy};
  return cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3(RTHREADS), params, smem, s);
}

template <int D, typename KVT>
cudaError_t launch_ring_g(int G, const MegaArgs& a, int grid, cudaStream_t s) {
  switch (G) {
    case 1: return launch_ring_t<D, 1, KVT>(a, grid, s);
    case 2: return launch_ring_t<D, 2, KVT>(a, grid, s);
    case 4: return launch_ring_t<D, 4, KVT>(a, grid, s);
    case 8: return launch_ring_t<D, 8, KVT>(a, grid, s);
  }
  return cudaErrorInvalidValue;
}

template <typename KVT>
cudaError_t launch_ring_d(const MegaArgs& a, int grid, cudaStream_t s) {
  const int G = a.q_heads / a.kv_heads;
  switch (a.head_dim) {
    case 32: return launch_ring_g<32, KVT>(G, a, grid, s);
    case 64: return launch_ring_g<64, KVT>(G, a, grid, s);
    case 128: return launch_ring_g<128, KVT>(G, a, grid, s);
  }
  return cudaErrorInvalidValue;
}

bool seg_ok(int K) { return K == SEG || K == 2 * SEG || K == 4 * SEG; }

}  // namespace

bool decode_ring_supported(int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim) {
  if (kv_heads < 1 || q_heads % kv_heads) return false;
  const int G = q_heads / kv_heads;
  if (!(G == 1 || G == 2 || G == 4 || G == 8)) return false;
  if (!(head_dim == 32 || head_dim == 64 || head_dim == 128)) return false;
  if (H != SEG || !seg_ok(QD) || !seg_ok(I)) return false;
  if (((QD + 2 * KD) / 2) % 4 || V % 8 || I % 64) return false;
  return ring_smem_bytes(G, head_dim) + 1024 <= 227 * 1024;
}

cudaError_t launch_decode_ring(const MegaArgs& a, int sm_count, cudaStream_t s) {
  if (!decode_ring_supported(a.H, a.QD, a.KD, a.I, a.V, a.q_heads, a.kv_heads, a.head_dim)) return cudaErrorInvalidValue;
  return a.kv_fp32 ? launch_ring_d<float>(a, sm_count, s) : launch_ring_d<uint16_t>(a, sm_count, s);
}

}  // namespace llmc
