// decode_ring_mma.cu -- the shared-memory-ring decode kernel (decode_ring.cu) with the slot arithmetic on the tensor
// cores: mma.sync.m16n8k16 with the WEIGHTS as the 8-column B operand and the activation vector as three rows of
// the A operand (x = hi + mid + lo, three bf16 values = 24 mantissa bits, so every product is exact and the sum is
// accumulated in fp32, as with the FMA loop).
//
// Why: in decode_ring.cu a warp spends ~1 000 cycles on its 8 KB share of a slot (270 instructions: 16 LDS.128,
// 128 unpack, 128 FMA, shuffles) -- latency that sits on the critical path of every short phase (QKV, O, the first
// slots after each barrier) and whose schedule ptxas decides (an unrelated change elsewhere in
// the kernel made it issue each load just before its use, +45 % per slot).  Here a slot is 8 rows x 2048 k; all 8
// consumer warps work on the SAME slot, warp w on k range [256 w, 256 w + 256): 8 LDS.128 + 16 MMAs per warp, the
// weights go to registers and the slot is handed back before the first MMA.  A lane's 16 bytes of a weight row
// serve two MMAs (the k order inside a 32-k block is permuted identically for A and B, which a dot product does not
// see).  Partial sums (3 A rows x 8 warps) meet through shuffles and one block barrier per row group.
//   * ring: 6 slots of 8 rows, rows 4160 bytes apart (64 bytes of padding make the fragment loads conflict-free), one
//     4 KB bulk copy per row segment;
//   * K > 2048 (O with 4096 columns, down with 8192): a row group is K / 2048 consecutive slots of the same block, the
//     accumulators stay in registers across them;
//   * the activation vector is split into its three bf16 parts once per block per phase (shared memory), each warp
//     then picks up its A fragments with one 16-byte load per 32 k.
// Everything else -- producer thread, grid barriers, attention phase, schedule rotation -- is decode_ring.cu's.
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

constexpr int MROW = SEG * 2 + 64;        // bytes between the 8 rows of a slot (padding: conflict-free fragment loads)
constexpr int MSLOT_BYTES = 8 * MROW;     // one ring slot: 8 row segments of SEG bf16
constexpr int MAXSEG = 4;                 // K <= 4 * SEG

__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  // A = 16 bytes of one weight row: (a.x, a.y) are rows 0..7 of the A tile, (a.z, a.w) rows 8..15 -- see slot_mma
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a.x), "r"(a.z), "r"(a.y), "r"(a.w), "r"(b0), "r"(b1));
}
// x -> its three bf16 parts (hi + mid + lo == x to 24 bits), packed in pairs
__device__ __forceinline__ void split3(float x0, float x1, uint32_t& h, uint32_t& m, uint32_t& l) {
  h = pack_bf16x2(x0, x1);
  const float r0 = x0 - bf16lo(h), r1 = x1 - bf16hi(h);
  m = pack_bf16x2(r0, r1);
  l = pack_bf16x2(r0 - bf16lo(m), r1 - bf16hi(m));
}

template <int D, int G, typename KVT>
__global__ void __launch_bounds__(RTHREADS, 1) decode_ring_mma_kernel(const MegaArgs a) {
  constexpr int PS = D + 2;
  using LM = LaneMap<D, G, KVT>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* att = reinterpret_cast<float*>(smem_raw + (size_t)NSLOT * MSLOT_BYTES);  // attention phase scratch / activation parts
  __shared__ __align__(8) unsigned long long s_bar[2 * NSLOT];                     // full[NSLOT], empty[NSLOT]
  __shared__ float s_acc[RW][8][8];                                                // [warp][row group of the batch][row]
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
      mbar_init(empty_bar(s), RW);   // every consumer warp takes its k range of every slot
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
      const int nseg = p.K / SEG, ngroups = p.ntasks / 4;   // a row group = 4 row pairs = 8 rows, nseg slots
      int first = (int)blockIdx.x - rot;
      if (first < 0) first += grid;
      auto issue = [&](int rg, int sg) {
        const unsigned r = seq % NSLOT, use = seq / NSLOT;
        ++seq;
        if (use > 0) {
          const long long t0 = stats ? clock64() : 0;
          mbar_wait(empty_bar(r), (use - 1) & 1);
          if (stats) t_wait += clock64() - t0;
        }
        mbar_expect_tx(full_bar(r), 8 * SEG * 2);
        const uint32_t dst = ring_s + r * MSLOT_BYTES;
#pragma unroll
        for (int pr = 0; pr < 4; ++pr) {   // slot rows 0..3: first row of each pair, rows 4..7: second row
          int r0, r1;
          pair_rows(p, rg * 4 + pr, r0, r1);
          bulk_g2s(dst + pr * MROW, p.W + (size_t)r0 * p.K + (size_t)sg * SEG, SEG * 2, full_bar(r));
          bulk_g2s(dst + (4 + pr) * MROW, p.W + (size_t)r1 * p.K + (size_t)sg * SEG, SEG * 2, full_bar(r));
        }
      };
      if (nseg == 1) {
        for (int rg = first; rg < ngroups; rg += grid) issue(rg, 0);
      } else {
        // K > SEG: segment-major over this block's (at most two) row groups, the order in which the consumers take them
        for (int sg = 0; sg < nseg; ++sg)
          for (int rg = first; rg < ngroups && rg < first + 2 * grid; rg += grid) issue(rg, sg);
      }
      rot = (rot + ngroups) % grid;
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
  const int fg = lane >> 2, ft = lane & 3;   // fragment coordinates: B column / A row fg, k slot ft
  unsigned seq = 0;
  int rot = 0;
  unsigned bar_target = gridDim.x;
  int tix = 0;
  // B fragments of the activation vector for this warp's k range [256 warp, +256) of the current 2048-k segment.  The
  // weights are the A operand: a lane's 16 bytes of weight row fg, k = 32 kb + 8 ft + 0..7, fill BOTH row halves of
  // the 16-row A tile -- rows 0..7 take the first four k (logical k slots 2ft, 2ft+1, 2ft+8, 2ft+9), rows 8..15 the
  // last four -- so one MMA covers 8 weight rows x 32 k.  The B columns carry the activation for both halves: columns
  // 0..2 = hi / mid / lo of the first four k, columns 3..5 = hi / mid / lo of the last four, columns 6, 7 zero.  The
  // useful quadrants of the result are rows 0..7 x columns 0..2 and rows 8..15 x columns 3..5.
  uint32_t breg[8][2];
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

  // Activation vector -> A fragments.  The block reads the vector from L2 ONCE per 2048-k segment (8 floats per
  // thread; RMSNorm gain folded, sum of squares reduced through shared memory), splits it into its bf16 parts in
  // shared memory, and every warp picks up its fragments: one 16-byte load per 32 k.
  uint16_t* xp = reinterpret_cast<uint16_t*>(att);   // [3][SEG] hi, mid, lo (free outside the attention phase)
  auto stage_parts = [&](const float4& v0, const float4& v1) {
    uint32_t h[4], m[4], l[4];
    split3(v0.x, v0.y, h[0], m[0], l[0]);
    split3(v0.z, v0.w, h[1], m[1], l[1]);
    split3(v1.x, v1.y, h[2], m[2], l[2]);
    split3(v1.z, v1.w, h[3], m[3], l[3]);
    *reinterpret_cast<uint4*>(xp + tid * 8) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(xp + SEG + tid * 8) = make_uint4(m[0], m[1], m[2], m[3]);
    *reinterpret_cast<uint4*>(xp + 2 * SEG + tid * 8) = make_uint4(l[0], l[1], l[2], l[3]);
  };
  auto pickup = [&]() {
    // column fg: part fg % 3 of k = 32 kb + 8 ft + (fg < 3 ? 0..3 : 4..7); two registers (k pairs) per 32-k block
    const int part = fg < 3 ? fg : fg - 3;
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) {
      uint2 f = make_uint2(0u, 0u);
      if (fg < 6) f = *reinterpret_cast<const uint2*>(xp + part * SEG + warp * 256 + kb * 32 + ft * 8 + (fg < 3 ? 0 : 4));
      breg[kb][0] = f.x; breg[kb][1] = f.y;
    }
  };
  // K == SEG.  gain != NULL: returns 1/rms.
  auto load_x_full = [&](const float* src, const uint16_t* emb_row, const uint16_t* gain, float* x_out) -> float {
    const long long tx0 = stats ? clock64() : 0;
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
      stage_parts(v0, v1);
    }
    cbar();
    pickup();
    float rs = 1.f;
    if (gain) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < RW; ++w) t += s_red[w];
      rs = rsqrtf(t / (float)SEG + a.eps);
    }
    if (stats) t_x += clock64() - tx0 + (long long)(breg[0][0] == 0x12345678u);   // keeps the loads inside the window
    return rs;
  };
  // the 8 MMAs of one slot (one per 32-k block of this warp's range).  Four independent accumulator chains (an MMA
  // waits ~25 cycles for the previous one on the same accumulator), summed at the end.
  auto slot_mma = [&](unsigned r, float (&c)[4]) {
    const uint32_t base = ring_s + r * MSLOT_BYTES + (uint32_t)(fg * MROW + (warp * 256 + ft * 8) * 2);
    uint4 w[8];
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) w[kb] = lds128(base + kb * 64);
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_bar(r));   // the warp's share of the slot is in registers: hand it back
    float ch[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) ch[i][j] = 0.f;
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) mma_bf16(ch[kb & 3], w[kb], breg[kb][0], breg[kb][1]);
#pragma unroll
    for (int j = 0; j < 4; ++j) c[j] += (ch[0][j] + ch[1][j]) + (ch[2][j] + ch[3][j]);
  };
  // Partial sums of a row group (3 A rows x 8 warps) meet in two stages: shuffles inside the warp, then shared memory.
  // The exchange is batched: up to NGB row groups share one pair of block barriers, and the 4 * NGB row pairs of a
  // batch are finished (summed over the 8 warps, epilogue) by all 8 warps.
  constexpr int NGB = 8;
  int pend = 0, pend_rg0 = 0;
  auto put_group = [&](float (&c)[4], int rg) {
    // lane (fg, ft) holds C[row fg][cols 2ft, 2ft+1] in c[0], c[1] and C[row fg + 8][same cols] in c[2], c[3]; weight
    // row fg's sum = C[fg][0] + C[fg][1] + C[fg][2] + C[fg+8][3] + C[fg+8][4] + C[fg+8][5]
    float v = ft == 0 ? c[0] + c[1] : ft == 1 ? c[0] + c[3] : ft == 2 ? c[2] + c[3] : 0.f;
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    if (ft == 0) s_acc[warp][pend][fg] = v;
    if (pend == 0) pend_rg0 = rg;
    ++pend;
  };
  auto flush_groups = [&](float rs, auto&& epilogue) {   // block-uniform
    if (pend == 0) return;
    cbar();
    for (int pi = warp; pi < 4 * pend; pi += RW) {   // pair pi: group pi / 4 (row group pend_rg0 + (pi / 4) * grid), pair pi % 4
      const int gi = pi >> 2, pr = pi & 3;
      float s0 = lane < RW ? s_acc[lane][gi][pr] : 0.f;
      float s1 = lane < RW ? s_acc[lane][gi][4 + pr] : 0.f;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      }
      s0 = __shfl_sync(0xffffffffu, s0, 0);
      s1 = __shfl_sync(0xffffffffu, s1, 0);
      epilogue((pend_rg0 + gi * grid) * 4 + pr, s0 * rs, s1 * rs);
    }
    cbar();   // s_acc may be rewritten
    pend = 0;
  };
  // one GEMV phase with K == SEG: this block's row groups; all 8 warps work on every slot, warp w on k range [256 w, +256)
  auto run_phase = [&](const RPhase& p, float rs, auto&& epilogue) {
    const int ngroups = p.ntasks / 4;
    int first = (int)blockIdx.x - rot;
    if (first < 0) first += grid;
    for (int rg = first; rg < ngroups; rg += grid) {
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      const unsigned r = seq % NSLOT, par = (seq / NSLOT) & 1u;
      ++seq;
      const long long tf0 = stats ? clock64() : 0;
      mbar_wait(full_bar(r), par);
      const long long tf1 = stats ? clock64() : 0;
      if (stats) t_full += tf1 - tf0;
      slot_mma(r, c);
      put_group(c, rg);
      if (pend == NGB) flush_groups(rs, epilogue);
      if (stats) t_slots += clock64() - tf1;
    }
    flush_groups(rs, epilogue);
    rot = (rot + ngroups) % grid;
  };
  // K = nseg * SEG > SEG (no RMSNorm on these inputs): segment-major.  The block requests all segments of the
  // activation at once, then for each segment stages its parts, picks its fragments up and runs that segment's slot of
  // each of its (at most two) row groups, whose accumulators stay in registers; the producer issues the slots in the
  // same order.
  auto run_phase_segs = [&](const RPhase& p, const float* src, auto&& epilogue) {
    const int nseg = p.K / SEG, ngroups = p.ntasks / 4;
    int first = (int)blockIdx.x - rot;
    if (first < 0) first += grid;
    const long long tx0 = stats ? clock64() : 0;
    float4 v[MAXSEG][2];
#pragma unroll
    for (int sg = 0; sg < MAXSEG; ++sg)
      if (sg < nseg) {
        v[sg][0] = __ldcg(reinterpret_cast<const float4*>(src + sg * SEG + tid * 8));
        v[sg][1] = __ldcg(reinterpret_cast<const float4*>(src + sg * SEG + tid * 8 + 4));
      }
    float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
    const bool two = first + grid < ngroups;
#pragma unroll
    for (int sg = 0; sg < MAXSEG; ++sg)
      if (sg < nseg) {
        if (sg > 0) cbar();          // everyone has picked the previous segment up
        stage_parts(v[sg][0], v[sg][1]);
        cbar();
        pickup();
        if (stats && sg == 0) t_x += clock64() - tx0;
        if (first < ngroups) {
          const unsigned r = seq % NSLOT, par = (seq / NSLOT) & 1u;
          ++seq;
          mbar_wait(full_bar(r), par);
          slot_mma(r, c0);
        }
        if (two) {
          const unsigned r = seq % NSLOT, par = (seq / NSLOT) & 1u;
          ++seq;
          mbar_wait(full_bar(r), par);
          slot_mma(r, c1);
        }
      }
    if (first < ngroups) put_group(c0, first);
    if (two) put_group(c1, first + grid);
    flush_groups(1.f, epilogue);
    rot = (rot + ngroups) % grid;
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
        if (p.K == SEG) {
          load_x_full(a.attn, nullptr, nullptr, nullptr);
          run_phase(p, 1.f, resid_epilogue);
        } else {
          run_phase_segs(p, a.attn, resid_epilogue);
        }
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
        if (p.K == SEG) {
          load_x_full(a.act, nullptr, nullptr, nullptr);
          run_phase(p, 1.f, resid_epilogue);
        } else {
          run_phase_segs(p, a.act, resid_epilogue);
        }
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
  // the same region stages the three bf16 parts of an activation segment outside the attention phase
  if (att < (size_t)3 * SEG / 2) att = (size_t)3 * SEG / 2;
  return (size_t)NSLOT * MSLOT_BYTES + att * sizeof(float);
}

template <int D, int G, typename KVT>
cudaError_t launch_ring_t(const MegaArgs& a, int grid, cudaStream_t s) {
  auto kern = decode_ring_mma_kernel<D, G, KVT>;
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

bool decode_ring_mma_supported(int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim) {
  if (kv_heads < 1 || q_heads % kv_heads) return false;
  const int G = q_heads / kv_heads;
  if (!(G == 1 || G == 2 || G == 4 || G == 8)) return false;
  if (!(head_dim == 32 || head_dim == 64 || head_dim == 128)) return false;
  if (H != SEG || !seg_ok(QD) || !seg_ok(I)) return false;
  if (((QD + 2 * KD) / 2) % 4 || V % 8 || I % 64) return false;
  return ring_smem_bytes(G, head_dim) + 1024 <= 227 * 1024;
}

cudaError_t launch_decode_ring_mma(const MegaArgs& a, int sm_count, cudaStream_t s) {
  if (!decode_ring_mma_supported(a.H, a.QD, a.KD, a.I, a.V, a.q_heads, a.kv_heads, a.head_dim)) return cudaErrorInvalidValue;
  if (a.H / 8 > 2 * sm_count) return cudaErrorInvalidValue;   // K > 2048 phases: at most two row groups of 8 output rows per block
  return a.kv_fp32 ? launch_ring_d<float>(a, sm_count, s) : launch_ring_d<uint16_t>(a, sm_count, s);
}

}  // namespace llmc
