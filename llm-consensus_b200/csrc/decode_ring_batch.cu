// decode_ring_batch.cu -- the shared-memory-ring decode kernel for 1..4 sequences at once (one persistent cooperative
// launch for all steps of a llmc_decode call, every weight byte streamed ONCE per step for the whole batch).  Also the
// default batch-1 kernel for 2 048-wide models (second sequence absent): faster than decode_ring.cu.
//
// Same skeleton as decode_ring_mma.cu: a producer thread fills a 6-slot ring (slot = 8 weight rows x 2048 k, 8 bulk
// copies), the 8 consumer warps all work on the same slot (warp w: k range [256 w, +256)), the slot arithmetic is
// mma.sync.m16n8k16 with the WEIGHTS as the A operand -- a lane's 16 bytes of weight row fg fill both row halves of the
// 16-row tile (rows 0..7: its first four k, rows 8..15: its last four), so one MMA covers 8 rows x 32 k.  What changes
// is the B operand: its 8 columns carry a PAIR of sequences, each as two bf16 parts (x = hi + mid, 16 mantissa bits,
// the same precision as the hi + lo operand pairs of the GEMM decode path) -- columns 0..3 = {seq0 hi, seq0 mid,
// seq1 hi, seq1 mid} of the first four k, columns 4..7 the same of the last four.  Four sequences = two MMAs per 32 k
// against the same weight registers.
//   * whoever PRODUCES an activation leaves it as finished bf16 hi / mid planes, so the 148 consumers of the next
//     phase load fragments (8 bytes per 32 k and lane) instead of converting fp32 each: the attention output, the
//     SwiGLU activation, and the residual stream -- the O and down projections update it by load-add-store (one
//     owner per element) and publish it pre-multiplied by the next RMSNorm's gain; its sum of squares is collected in
//     64-bit fixed-point accumulators (integer addition: deterministic whatever the arrival order), 1/rms is applied
//     in the epilogue (it is a scalar factor).  Only layer 0 converts: its input is the embedding row;
//   * activation fragments go from L2 straight to registers -- no staging in shared memory, no block barrier;
//   * K > 2048 (down projection): segment-major as in decode_ring_mma.cu, the next segment's activation loads are in
//     flight while the current segment's slots run;
//   * partial sums (4 lanes x 8 warps per row and sequence) meet through shuffles and a batched shared-memory exchange;
//     the 32 lanes of a finishing warp are four groups of 8, one (row pair, sequence) each;
//   * attention: one block per (sequence, kv head, split); sequences have their own lengths and pages;
//   * greedy argmax per sequence, tokens of all steps left in out_tokens[seq * out_stride + step].
// HBM-bound; algorithmic bytes per step: the weights once + each sequence's K/V cache.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary).
#include <cstdlib>
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
constexpr int NGB = 4;                    // row groups per exchange batch

__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  // A = 16 bytes of one weight row: (a.x, a.y) are rows 0..7 of the A tile, (a.z, a.w) rows 8..15
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a.x), "r"(a.z), "r"(a.y), "r"(a.w), "r"(b0), "r"(b1));
}

// two floats -> packed bf16 pair in ONE instruction (F2FP.BF16.F32.PACK_AB; two scalar converts + a merge are several
// times slower, and a warp converts 64..128 values per activation segment here)
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// NP = pairs of sequences (1: batch 2; 2: batch 3 or 4).  ONE (with NP = 1): a single sequence whose activation takes
// three of the four columns of a half -- hi, mid, lo: 24 mantissa bits, every product exact as in the FMA kernels.
template <int D, int G, typename KVT, int NP, bool ONE = false>
__global__ void __launch_bounds__(RTHREADS, 1) decode_ring_batch_kernel(const MegaArgs a) {
  static_assert(!ONE || NP == 1, "the single-sequence variant uses one column group");
  constexpr int NB = 2 * NP;
  constexpr int PS = D + 2;
  constexpr int hd2 = D / 2;
  using LM = LaneMap<D, G, KVT>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* att = reinterpret_cast<float*>(smem_raw + (size_t)NSLOT * MSLOT_BYTES);  // attention phase scratch
  __shared__ __align__(8) unsigned long long s_bar[2 * NSLOT];                     // full[NSLOT], empty[NSLOT]
  __shared__ float s_acc[RW][NGB][NB][8];   // [warp][row group of the batch][sequence][row]
  __shared__ float s_rope[NB][2][hd2];      // this step's RoPE rows (cos, sin) of each sequence
  __shared__ float s_red[RW][NB];           // sum of squares per warp and sequence
  __shared__ float s_ss[RW][NB];            // ... of the rows a residual phase finished in this block
  __shared__ float s_val[RW][NB];
  __shared__ int s_idx[RW][NB];
  __shared__ int s_tok[NB], s_slot[NB], s_pos0[NB];
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int grid = (int)gridDim.x;
  const int B = a.B;
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
  if (tid < NB) {
    const int sl = tid < B ? a.slots[tid] : 0;
    s_slot[tid] = sl;
    s_pos0[tid] = tid < B ? a.seq_len[sl] : 0;
    s_tok[tid] = tid < B ? a.next_token[sl] : 0;
  }
  __syncthreads();

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

  // ======================================================================== producer (as in decode_ring_mma.cu)
  if (warp >= RW) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    if (warp != RW || lane != 0) return;
    unsigned seq = 0;
    int rot = 0;
    auto produce = [&](const RPhase& p) {
      const int nseg = p.K / SEG, ngroups = p.ntasks / 4;   // a row group = 4 row pairs = 8 rows, nseg slots
      int first = (int)blockIdx.x - rot;
      if (first < 0) first += grid;
      auto issue = [&](int rg, int sg) {
        const unsigned r = seq % NSLOT, use = seq / NSLOT;
        ++seq;
        if (use > 0) mbar_wait(empty_bar(r), (use - 1) & 1);
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
    return;
  }

  // ======================================================================== consumers
  asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
  // fragment coordinates: B column / A row fg, k slot ft.  Column fg = 4 half + 2 sl + part: half = which four of the
  // lane's eight k, sl = sequence of the pair, part = hi / mid.
  const int fg = lane >> 2, ft = lane & 3;
  // (ONE: column = 4 half + part, part 3 unused)
  const int half = fg >> 2, sl = ONE ? 0 : (fg >> 1) & 1, part = ONE ? (fg & 3) : (fg & 1);
  const bool col_live = !ONE || part < 3;
  const int klane = warp * 256 + ft * 8 + half * 4;   // + 32 kb: the four k this lane supplies per 32-k block
  // the finishing side: a warp's 32 lanes are four groups of 8 (one lane per consumer warp's partial sum); group j works
  // for sequence myseq
  const int wl = lane & 7, lj = lane >> 3;
  const int myseq = NB == 4 ? lj : (lj & 1);
  const int pi_off = NB == 4 ? 0 : (lj >> 1);
  constexpr int PPP = NB == 4 ? 1 : 2;   // row pairs a warp finishes per pass
  unsigned seq = 0;
  int rot = 0;
  unsigned bar_target = gridDim.x;
  uint32_t breg[NP][8][2];   // B fragments of the current 2048-k segment, this warp's k range
  const float scale = rsqrtf((float)D);
  // blocks per (sequence, kv head) the attention phase can use
  const int nspb = max(1, min(RING_MAX_SPLIT, grid / (B * a.kv_heads)));

  int tix = 0;
#ifdef LLMC_RINGB_STATS   // cycle counters of block 0 / warp 0 under LLMC_MEGA_TRACE: compiled out by default -- ten more live
                          // registers through the whole kernel are not free here (build with -DLLMC_RINGB_STATS to get them)
  const bool stats = a.trace && blockIdx.x == 0 && warp == 0;
#else
  constexpr bool stats = false;
#endif
  long long t_full = 0, t_x = 0, t_bar = 0, t_slots = 0, t_flush = 0;
  const long long t_begin = clock64();
  auto grid_barrier = [&]() {
    const long long tb0 = stats ? clock64() : 0;
    cbar();
    if (tid == 0) {
      if (a.trace && blockIdx.x == 0) a.trace[2 * tix] = gtime();   // LLMC_MEGA_TRACE: block 0's (arrive, open) times
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.barrier) : "memory");
      unsigned v;
      do {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.barrier) : "memory");
      } while (v < bar_target);
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.barrier) : "memory");
      if (a.trace && blockIdx.x == 0) a.trace[2 * tix + 1] = gtime();
    }
    cbar();
    if (stats) t_bar += clock64() - tb0;
    bar_target += gridDim.x;
    ++tix;
  };

  // ---- activation fragments.  raw loads (one 16-byte L2 load per 32-k block and pair) and their conversion are
  // separate, so a caller can keep the next segment's loads in flight behind the current segment's slots.
  auto load_raw = [&](float4 (&raw)[NP][8], const float* src, size_t stride, int koff) {
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int s = 2 * p + sl;
#pragma unroll
      for (int kb = 0; kb < 8; ++kb)
        raw[p][kb] = (s < B && col_live) ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)s * stride + koff + klane + kb * 32))
                           : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  auto to_frag = [&](const float4& v, uint32_t& b0, uint32_t& b1) {
    uint32_t h0 = pack2(v.x, v.y), h1 = pack2(v.z, v.w);
    if (part) {
      float r0 = v.x - bf16lo(h0), r1 = v.y - bf16hi(h0), r2 = v.z - bf16lo(h1), r3 = v.w - bf16hi(h1);
      h0 = pack2(r0, r1);
      h1 = pack2(r2, r3);
      if constexpr (ONE) {
        if (part == 2) {
          h0 = pack2(r0 - bf16lo(h0), r1 - bf16hi(h0));
          h1 = pack2(r2 - bf16lo(h1), r3 - bf16hi(h1));
        }
      }
    }
    b0 = h0;
    b1 = h1;
  };
  auto convert = [&](const float4 (&raw)[NP][8]) {
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
      for (int kb = 0; kb < 8; ++kb) to_frag(raw[p][kb], breg[p][kb][0], breg[p][kb][1]);
  };
  // inputs that the previous phase left as bf16 planes (attention output, SwiGLU activation): this lane's part of its
  // four k per 32-k block is one 8-byte load, already in fragment form
  auto load_planes_raw = [&](uint2 (&raw)[NP][8], const uint16_t* hi, const uint16_t* lo, const uint16_t* lo2, size_t stride, int koff) {
    const uint16_t* plane = part == 0 ? hi : part == 1 ? lo : lo2;
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int s = 2 * p + sl;
#pragma unroll
      for (int kb = 0; kb < 8; ++kb)
        raw[p][kb] = (s < B && col_live) ? __ldcg(reinterpret_cast<const uint2*>(plane + (size_t)s * stride + koff + klane + kb * 32))
                           : make_uint2(0u, 0u);
    }
  };
  auto take_planes = [&](const uint2 (&raw)[NP][8]) {
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
      for (int kb = 0; kb < 8; ++kb) {
        breg[p][kb][0] = raw[p][kb].x;
        breg[p][kb][1] = raw[p][kb].y;
      }
  };
  auto load_planes = [&](const uint16_t* hi, const uint16_t* lo, const uint16_t* lo2, size_t stride) {
    const long long tx0 = stats ? clock64() : 0;
    uint2 raw[NP][8];
    load_planes_raw(raw, hi, lo, lo2, stride, 0);
    take_planes(raw);
    if (stats) t_x += clock64() - tx0 + (long long)(breg[0][0][0] == 0x12345678u);
  };
  // K == SEG input with optional RMSNorm gain (sum of squares -> s_red, read by the epilogue side) and, in layer 0, the
  // embedding rows as the source (block 0 also writes them out as the residual stream)
  auto load_x = [&](const float* src, size_t stride, const uint16_t* gain, bool embed, bool write_x) {
    const long long tx0 = stats ? clock64() : 0;
    float4 raw[NP][8];
    if (embed) {
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        const int s = 2 * p + sl;
        const uint16_t* er = a.emb + (size_t)s_tok[s] * H;
#pragma unroll
        for (int kb = 0; kb < 8; ++kb) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (s < B && col_live) {
            const uint2 e = *reinterpret_cast<const uint2*>(er + klane + kb * 32);
            v = make_float4(bf16lo(e.x), bf16hi(e.x), bf16lo(e.y), bf16hi(e.y));
            if (write_x && part == 0) *reinterpret_cast<float4*>(a.x + (size_t)s * H + klane + kb * 32) = v;
          }
          raw[p][kb] = v;
        }
      }
    } else {
      load_raw(raw, src, stride, 0);
    }
    if (gain) {
      uint2 gb[8];
#pragma unroll
      for (int kb = 0; kb < 8; ++kb) gb[kb] = *reinterpret_cast<const uint2*>(gain + klane + kb * 32);
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        float ss = 0.f;
#pragma unroll
        for (int kb = 0; kb < 8; ++kb) {
          float4& v = raw[p][kb];
          ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
          v.x *= bf16lo(gb[kb].x); v.y *= bf16hi(gb[kb].x); v.z *= bf16lo(gb[kb].y); v.w *= bf16hi(gb[kb].y);
        }
        // the hi and mid lanes of a column pair loaded the same values: count them once.  Sum over the k slots (lane
        // bits 0, 1) and the two halves (lane bit 4).
        if (part) ss = 0.f;
        ss += __shfl_xor_sync(0xffffffffu, ss, 1);
        ss += __shfl_xor_sync(0xffffffffu, ss, 2);
        ss += __shfl_xor_sync(0xffffffffu, ss, 16);
        if (ft == 0 && half == 0 && part == 0) s_red[warp][2 * p + sl] = ss;
      }
    }
    convert(raw);
    if (stats) t_x += clock64() - tx0 + (long long)(breg[0][0][0] == 0x12345678u);   // keeps the loads inside the window
  };

  // the MMAs of one slot: 8 LDS.128, hand the slot back, then one MMA per 32-k block and pair (four accumulator
  // chains in all)
  constexpr int NCH = 4 / NP;
  auto slot_mma = [&](unsigned r, float (&c)[NP][4]) {
    const uint32_t base = ring_s + r * MSLOT_BYTES + (uint32_t)(fg * MROW + (warp * 256 + ft * 8) * 2);
    uint4 w[8];
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) w[kb] = lds128(base + kb * 64);
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_bar(r));   // the warp's share of the slot is in registers: hand it back
    float ch[NP][NCH][4];
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
      for (int i = 0; i < NCH; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) ch[p][i][j] = 0.f;
#pragma unroll
    for (int kb = 0; kb < 8; ++kb)
#pragma unroll
      for (int p = 0; p < NP; ++p) mma_bf16(ch[p][kb % NCH], w[kb], breg[p][kb][0], breg[p][kb][1]);
#pragma unroll
    for (int p = 0; p < NP; ++p)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < NCH; ++i) t += ch[p][i][j];
        c[p][j] += t;
      }
  };
  // lane (fg, ft) holds C[row fg][cols 2ft, 2ft+1] in c[0], c[1] and C[row fg + 8][same cols] in c[2], c[3].  Useful:
  // rows 0..7 x columns 0..3 and rows 8..15 x columns 4..7; sequence sl of the pair: columns 2 sl, 2 sl + 1 (+ 4).
  float rs_given = 1.f;       // 1/rms of sequence myseq when the input came as pre-scaled planes (load_xn)
  bool rs_is_given = false;
  int pend = 0, pend_rg0 = 0;
  auto put_group = [&](float (&c)[NP][4], int rg) {
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      float v = ft < 2 ? c[p][0] + c[p][1] : c[p][2] + c[p][3];
      v += __shfl_xor_sync(0xffffffffu, v, 2);   // first-half + second-half columns of the same sequence
      if constexpr (ONE) {
        v += __shfl_xor_sync(0xffffffffu, v, 1);   // (hi, mid) + (lo, unused)
        if (ft == 0) s_acc[warp][pend][0][fg] = v;
      } else {
        if (ft < 2) s_acc[warp][pend][2 * p + ft][fg] = v;
      }
    }
    if (pend == 0) pend_rg0 = rg;
    ++pend;
  };
  // epilogue(seq, task, v0, v1): called by the 8 lanes of a group with the finished pair of rows (2 task, 2 task + 1 in
  // pair_rows order) of one sequence; lane wl == 0 of the group stores
  auto flush_groups = [&](bool norm, int K, auto&& epilogue, bool last = false) {   // block-uniform
    if (pend == 0) return;
    const long long tf0 = stats ? clock64() : 0;
    cbar();
    float rs = 1.f;
    if (norm) {
      if (rs_is_given) {
        rs = rs_given;
      } else {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < RW; ++w) t += s_red[w][myseq];
        rs = rsqrtf(t / (float)K + a.eps);
      }
    }
    for (int pi0 = warp * PPP; pi0 < 4 * pend; pi0 += RW * PPP) {
      const int pi = pi0 + pi_off;   // (4 * pend is a multiple of PPP: never past the end)
      const int gi = pi >> 2, pr = pi & 3;
      float s0 = s_acc[wl][gi][myseq][pr];
      float s1 = s_acc[wl][gi][myseq][4 + pr];
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      }
      if (myseq < B) epilogue(myseq, (pend_rg0 + gi * grid) * 4 + pr, s0 * rs, s1 * rs);
    }
    // s_acc may be rewritten.  After a phase's last exchange the next block barrier (grid barrier, residual finish, argmax)
    // does that: skipped there in the one-pair kernel (batch 2: - 1.4 %; the two-pair kernel measured 3 % slower with it)
    if (!(last && NP == 1)) cbar();
    if (stats) t_flush += clock64() - tf0;
    pend = 0;
  };
  // one GEMV phase with K == SEG (fragments already in breg)
  auto run_phase = [&](const RPhase& p, bool norm, auto&& epilogue) {
    const int ngroups = p.ntasks / 4;
    int first = (int)blockIdx.x - rot;
    if (first < 0) first += grid;
    for (int rg = first; rg < ngroups; rg += grid) {
      float c[NP][4];
#pragma unroll
      for (int p2 = 0; p2 < NP; ++p2)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[p2][j] = 0.f;
      const unsigned r = seq % NSLOT, par = (seq / NSLOT) & 1u;
      ++seq;
      const long long tw0 = stats ? clock64() : 0;
      mbar_wait(full_bar(r), par);
      const long long tw1 = stats ? clock64() : 0;
      slot_mma(r, c);
      put_group(c, rg);
      if (stats) { t_full += tw1 - tw0; t_slots += clock64() - tw1; }
      if (pend == NGB) flush_groups(norm, p.K, epilogue);
    }
    flush_groups(norm, p.K, epilogue, true);
    rot = (rot + ngroups) % grid;
  };
  // K = nseg * SEG > SEG: segment-major over this block's (at most two) row groups, accumulators in registers; the next
  // segment's activation loads fly while this segment's slots run
  auto run_phase_segs = [&](const RPhase& p, const uint16_t* hi, const uint16_t* lo, const uint16_t* lo2, size_t stride, auto&& epilogue) {
    const int nseg = p.K / SEG, ngroups = p.ntasks / 4;
    int first = (int)blockIdx.x - rot;
    if (first < 0) first += grid;
    uint2 raw[NP][8];
    load_planes_raw(raw, hi, lo, lo2, stride, 0);
    float c0[NP][4], c1[NP][4];
#pragma unroll
    for (int p2 = 0; p2 < NP; ++p2)
#pragma unroll
      for (int j = 0; j < 4; ++j) c0[p2][j] = c1[p2][j] = 0.f;
    const bool two = first + grid < ngroups;
#pragma unroll
    for (int sg = 0; sg < MAXSEG; ++sg)
      if (sg < nseg) {
        take_planes(raw);
        if (sg + 1 < nseg) load_planes_raw(raw, hi, lo, lo2, stride, (sg + 1) * SEG);
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
    flush_groups(false, p.K, epilogue, true);
    rot = (rot + ngroups) % grid;
  };
  // ---- residual phases (O and down projection): x += W y.  Every row pair has ONE finishing lane group per sequence,
  // so the update is load - add - store, with the old value requested at the start of the phase.  The same lane then
  // leaves the new value, multiplied by the gain of the RMSNorm that reads it next, as bf16 hi / mid planes -- the
  // form in which the next phase's fragments are loaded -- and adds its square to this block's share of the sum.
  constexpr bool XN = !ONE;   // (the three-part variant keeps converting the fp32 stream itself)
  float2 xo = make_float2(0.f, 0.f);
  uint32_t gpair = 0u;
  float ssl = 0.f;
  int rphase = 0;   // residual phases done so far in this launch
  auto resid_prefetch = [&](const RPhase& p, const uint16_t* gain_next) {
    if constexpr (XN) {
      const int ngroups = p.ntasks / 4;
      int first = (int)blockIdx.x - rot;
      if (first < 0) first += grid;
      const int mypend = (first < ngroups ? 1 : 0) + (first + grid < ngroups ? 1 : 0);
      const int pi = warp * PPP + pi_off;   // the one exchange of the phase: this lane group's row pair (flush_groups)
      ssl = 0.f;
      if (pi < 4 * mypend && myseq < B && wl == 0) {
        const int task = (first + (pi >> 2) * grid) * 4 + (pi & 3);
        xo = __ldcg(reinterpret_cast<const float2*>(a.x + (size_t)myseq * H + 2 * task));
        gpair = *reinterpret_cast<const uint32_t*>(gain_next + 2 * task);
      }
    }
  };
  auto resid_epilogue = [&](int s, int task, float v0, float v1) {
    if (wl == 0) {
      if constexpr (XN) {
        const float x0 = xo.x + v0, x1 = xo.y + v1;
        *reinterpret_cast<float2*>(a.x + (size_t)s * H + 2 * task) = make_float2(x0, x1);
        ssl = x0 * x0 + x1 * x1;
        const float y0 = x0 * bf16lo(gpair), y1 = x1 * bf16hi(gpair);
        const uint32_t h = pack2(y0, y1);
        *reinterpret_cast<uint32_t*>(a.xn_hi + (size_t)s * H + 2 * task) = h;
        *reinterpret_cast<uint32_t*>(a.xn_lo + (size_t)s * H + 2 * task) = pack2(y0 - bf16lo(h), y1 - bf16hi(h));
      } else {
        // one writer per row and sequence: a fire-and-forget reduction (RED) gives the same fp32 sum as load-add-store
        atomicAdd(a.x + (size_t)s * H + 2 * task, v0);
        atomicAdd(a.x + (size_t)s * H + 2 * task + 1, v1);
      }
    }
  };
  // this block's share of sum(x^2) per sequence -> ss_part[seq][block] (zero when it finished no row)
  auto resid_finish = [&]() {
    if constexpr (XN) {
      float t = ssl;
      if constexpr (NB == 2) t += __shfl_xor_sync(0xffffffffu, t, 16);   // lane groups j and j + 2: the same sequence
      if (wl == 0 && lj < NB) s_ss[warp][myseq] = t;
      cbar();
      // The blocks' shares meet in a 64-bit fixed-point accumulator (integer addition is associative, so the total does
      // not depend on the order of arrival: deterministic) -- one 8-byte word per sequence for every reader, where 148
      // per-block floats would be a hot spot in L2 (measured: + 2 us per reading phase).  Four slots used in turn;
      // block 0 clears the one that is written two residual phases from now.
      if (tid < NB) {
        float u = 0.f;
#pragma unroll
        for (int w = 0; w < RW; ++w) u += s_ss[w][tid];
        atomicAdd(a.ss_acc + (rphase & 3) * 4 + tid, __float2ull_rn(u * 16777216.0f));
        if (blockIdx.x == 0) a.ss_acc[((rphase + 2) & 3) * 4 + tid] = 0ull;
      }
      ssl = 0.f;
      ++rphase;
    }
  };
  // RMSNorm input of the phases after a residual phase: finished fragments + 1/rms from the blocks' shares
  auto load_xn = [&]() {
    load_planes(a.xn_hi, a.xn_lo, nullptr, (size_t)H);
    const unsigned long long acc = myseq < B ? __ldcg(a.ss_acc + ((rphase - 1) & 3) * 4 + myseq) : 0ull;
    const float t = (float)acc * (1.0f / 16777216.0f);
    rs_given = rsqrtf(t / (float)H + a.eps);
    rs_is_given = true;
  };

  for (int step = 0; step < a.nsteps; ++step) {
    // this step's RoPE rows
    for (int i = tid; i < NB * hd2; i += RW * 32) {
      const int s = i / hd2, j = i - s * hd2;
      const int pos = s_pos0[s] + step;
      s_rope[s][0][j] = s < B ? a.rope_cos[(size_t)pos * hd2 + j] : 1.f;
      s_rope[s][1][j] = s < B ? a.rope_sin[(size_t)pos * hd2 + j] : 0.f;
    }
    // (visible to the QKV epilogue: flush_groups starts with a block barrier)

    // the attention task of this block: (sequence, kv head, split)
    const int at_kvh = blockIdx.x % a.kv_heads, at_t = blockIdx.x / a.kv_heads;
    const int at_s = at_t % B, at_split = at_t / B;
    const int at_pos = s_pos0[at_s] + step, at_ctx = at_pos + 1;
    const int at_ntiles = (at_ctx + TP - 1) / TP;
    const int at_nsp = max(1, min(min(nspb, a.max_tiles), (at_ntiles + a.attn_tiles_per_block - 1) / a.attn_tiles_per_block));
    const bool at_active = at_split < at_nsp && at_split < nspb;
    const int* at_bt = a.block_table + (size_t)s_slot[at_s] * a.bt_stride;

    for (int l = 0; l < a.L; ++l) {
      const MegaLayer& ly = a.layers[l];
      KVT* kpool = reinterpret_cast<KVT*>(ly.kp);
      KVT* vpool = reinterpret_cast<KVT*>(ly.vp);

      // ---------------- P1: RMSNorm -> QKV -> RoPE -> KV append
      {
        if (XN && l > 0) {
          load_xn();
        } else {
          rs_is_given = false;
          load_x(a.x, (size_t)H, ly.g1, l == 0, l == 0 && blockIdx.x == 0);
        }
        run_phase(phase_desc(l, 0), true, [&](int s, int g, float v0, float v1) {
          const int h = g / hd2, i = g - h * hd2;
          if (h < a.q_heads + a.kv_heads) {
            const float c = s_rope[s][0][i], sn = s_rope[s][1][i];
            const float o0 = v0 * c - v1 * sn, o1 = v1 * c + v0 * sn;
            v0 = o0; v1 = o1;
          }
          if (wl == 0) {
            if (h < a.q_heads) {
              a.q[(size_t)s * QD + h * D + i] = v0;
              a.q[(size_t)s * QD + h * D + i + hd2] = v1;
            } else {
              const bool isk = h < a.q_heads + a.kv_heads;
              const int kvh = isk ? h - a.q_heads : h - a.q_heads - a.kv_heads;
              const int pos = s_pos0[s] + step;
              const int page = a.block_table[(size_t)s_slot[s] * a.bt_stride + pos / a.page_size];
              KVT* dst = (isk ? kpool : vpool) + (((size_t)page * a.kv_heads + kvh) * a.page_size + pos % a.page_size) * D + i;
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
        // ask L2 for the K/V tiles this block will read in the attention phase (written by earlier steps, long evicted
        // by the weight stream)
        if (at_active) {
          constexpr int TILE_LINES = TP * D * (int)sizeof(KVT) / 128;
          for (int tile = at_split + at_nsp * warp; tile < at_ntiles; tile += at_nsp * RW) {
            const int p0 = tile * TP;
            const int page = at_bt[p0 / a.page_size];
            const size_t base = (((size_t)page * a.kv_heads + at_kvh) * a.page_size + (p0 % a.page_size)) * D;
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

      // ---------------- P2: attention: block -> (sequence, kv head, split)
      {
        float* qs = att;                                           // [G][D]
        float* part_rec = att + G * D;                             // [RW][G][PS]
        float* pw = part_rec + (size_t)RW * G * PS + (size_t)warp * G * (TP + 1);
        if (at_active) {
          const int kvh = at_kvh, split = at_split, nsp = at_nsp;
          for (int i = tid; i < G * D; i += RW * 32) qs[i] = __ldcg(a.q + (size_t)at_s * QD + (size_t)kvh * G * D + i);
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
          for (int tile = split + nsp * warp; tile < at_ntiles; tile += nsp * RW) {
            const int p0 = tile * TP;
            const int valid = min(TP, at_ctx - p0);
            const int page = at_bt[p0 / a.page_size];
            const size_t base = (((size_t)page * a.kv_heads + kvh) * a.page_size + (p0 % a.page_size)) * D;
            attn_tile_step<D, G, KVT, true, 16>(qs, pw, kpool + base + c * LM::EPV, vpool + base + c * LM::EPV, valid,
                                                scale, m, lsum, o);
          }
          attn_write_record<D, G, KVT>(part_rec + (size_t)warp * G * PS, m, lsum, o);
          cbar();
          // the result goes out as bf16 hi / mid planes, the form the O projection's fragments take
          auto put_out = [&](int idx, const float (&v)[D / 32]) {
            constexpr int DPL = D / 32;
#pragma unroll
            for (int i = 0; i < DPL; i += 2) {
              const uint32_t h = pack2(v[i], v[i + 1]);
              const float r0 = v[i] - bf16lo(h), r1 = v[i + 1] - bf16hi(h);
              const uint32_t m = pack2(r0, r1);
              *reinterpret_cast<uint32_t*>(a.attn_hi + (size_t)at_s * QD + idx + i) = h;
              *reinterpret_cast<uint32_t*>(a.attn_lo + (size_t)at_s * QD + idx + i) = m;
              if constexpr (ONE)
                *reinterpret_cast<uint32_t*>(a.attn_lo2 + (size_t)at_s * QD + idx + i) = pack2(r0 - bf16lo(m), r1 - bf16hi(m));
            }
          };
          float* scratch = a.attn_scratch + (size_t)(at_s * a.kv_heads + kvh) * nspb * G * PS;
          if (warp < G) {  // warp h merges head h over the block's warps
            constexpr int DPL = D / 32;
            const int h = warp;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < RW; ++w) M = fmaxf(M, part_rec[((size_t)w * G + h) * PS + D]);
            float L = 0.f, O[DPL];
#pragma unroll
            for (int i = 0; i < DPL; ++i) O[i] = 0.f;
            if (M != -INFINITY) {
#pragma unroll
              for (int w = 0; w < RW; ++w) {
                const float* pp = part_rec + ((size_t)w * G + h) * PS;
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
              for (int i = 0; i < DPL; ++i) O[i] *= inv;
              put_out((kvh * G + h) * D + lane * DPL, O);
            } else {
              float* rec = scratch + ((size_t)split * G + h) * PS;
#pragma unroll
              for (int i = 0; i < DPL; ++i) rec[lane * DPL + i] = O[i];
              if (lane == 0) {
                rec[D] = M;
                rec[D + 1] = L;
              }
            }
          }
          // the last block to finish this (sequence, kv head) merges its nsp split records
          if (nsp > 1) {
            __threadfence();
            cbar();
            if (tid == 0) s_last = (atomicAdd(a.attn_counters + at_s * a.kv_heads + kvh, 1u) == (unsigned)(nsp - 1)) ? 1 : 0;
            cbar();
          }
          if (nsp > 1 && s_last && warp < G) {
            __threadfence();
            constexpr int DPL = D / 32;
            const int h = warp;
            const float* rb = scratch + (size_t)h * PS;
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
              for (int i = 0; i < DPL; ++i) O[i] *= inv;
              put_out((kvh * G + h) * D + lane * DPL, O);
            };
            if (nsp <= 4) merge(std::integral_constant<int, 4>{});
            else if (nsp <= 8) merge(std::integral_constant<int, 8>{});
            else merge(std::integral_constant<int, RING_MAX_SPLIT>{});
            if (lane == 0 && h == 0) a.attn_counters[at_s * a.kv_heads + kvh] = 0u;
          }
        }
        grid_barrier();
      }

      // ---------------- P3: x += Wo attn
      {
        const RPhase p = phase_desc(l, 1);
        resid_prefetch(p, ly.g2);
        if (p.K == SEG) {
          load_planes(a.attn_hi, a.attn_lo, a.attn_lo2, (size_t)QD);
          run_phase(p, false, resid_epilogue);
        } else {
          run_phase_segs(p, a.attn_hi, a.attn_lo, a.attn_lo2, (size_t)QD, resid_epilogue);
        }
        resid_finish();
        grid_barrier();
      }

      // ---------------- P4: act = silu(Wg h) * (Wu h), h = RMSNorm(x)
      {
        if constexpr (XN) {
          load_xn();
        } else {
          rs_is_given = false;
          load_x(a.x, (size_t)H, ly.g2, false, false);
        }
        run_phase(phase_desc(l, 2), true, [&](int s, int g, float gt, float up) {
          if (wl == 0) {
            const float v = gt / (1.0f + __expf(-gt)) * up;
            const uint16_t h = f32_to_bf16_bits(v);
            const float r = v - __uint_as_float((uint32_t)h << 16);
            const uint16_t m = f32_to_bf16_bits(r);
            a.act_hi[(size_t)s * I + g] = h;
            a.act_lo[(size_t)s * I + g] = m;
            if constexpr (ONE) a.act_lo2[(size_t)s * I + g] = f32_to_bf16_bits(r - __uint_as_float((uint32_t)m << 16));
          }
        });
        grid_barrier();
      }

      // ---------------- P5: x += Wd act
      {
        const RPhase p = phase_desc(l, 3);
        resid_prefetch(p, l + 1 < a.L ? a.layers[l + 1].g1 : a.gf);
        if (p.K == SEG) {
          load_planes(a.act_hi, a.act_lo, a.act_lo2, (size_t)I);
          run_phase(p, false, resid_epilogue);
        } else {
          run_phase_segs(p, a.act_hi, a.act_lo, a.act_lo2, (size_t)I, resid_epilogue);
        }
        resid_finish();
        grid_barrier();
      }
    }

    // ---------------- LM head + greedy argmax per sequence (lowest index wins ties)
    {
      if constexpr (XN) {
        load_xn();
      } else {
        rs_is_given = false;
        load_x(a.x, (size_t)H, a.gf, false, false);
      }
      float best = -INFINITY;   // of sequence myseq, over the row pairs this lane group finished
      int bi = 0x7fffffff;
      const bool store_logits = a.logits != nullptr && step == a.nsteps - 1;
      run_phase(phase_desc(-1, 0), true, [&](int s, int g, float v0, float v1) {
        if (store_logits && wl == 0) *reinterpret_cast<float2*>(a.logits + (size_t)s * V + 2 * g) = make_float2(v0, v1);
        if (v0 > best || (v0 == best && 2 * g < bi)) { best = v0; bi = 2 * g; }
        if (v1 > best || (v1 == best && 2 * g + 1 < bi)) { best = v1; bi = 2 * g + 1; }
      });
      if constexpr (NB == 2) {   // lane groups j and j + 2 worked for the same sequence
        const float ov = __shfl_xor_sync(0xffffffffu, best, 16);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, 16);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
      }
      if (wl == 0 && lj < NB) {
        s_val[warp][myseq] = best;
        s_idx[warp][myseq] = bi;
      }
      cbar();
      if (tid < NB) {
        float bv = s_val[0][tid];
        int bx = s_idx[0][tid];
#pragma unroll
        for (int w = 1; w < RW; ++w)
          if (s_val[w][tid] > bv || (s_val[w][tid] == bv && s_idx[w][tid] < bx)) { bv = s_val[w][tid]; bx = s_idx[w][tid]; }
        a.part_val[(size_t)tid * grid + blockIdx.x] = bv;
        a.part_idx[(size_t)tid * grid + blockIdx.x] = bx;
      }
      grid_barrier();
      // every block reduces the per-block candidates itself: no second exchange.  Warp s takes sequence s.
      if (warp < B) {
        float bv = -INFINITY;
        int bx = 0x7fffffff;
        for (int j = lane; j < grid; j += 32) {
          const float v = __ldcg(a.part_val + (size_t)warp * grid + j);
          const int ix = __ldcg(a.part_idx + (size_t)warp * grid + j);
          if (v > bv || (v == bv && ix < bx)) { bv = v; bx = ix; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bx, o);
          if (ov > bv || (ov == bv && oi < bx)) { bv = ov; bx = oi; }
        }
        if (lane == 0) {
          s_tok[warp] = bx;
          if (blockIdx.x == 0) a.out_tokens[(size_t)warp * a.out_stride + step] = bx;
        }
      }
      cbar();
    }
  }
  if (stats && lane == 0) {
    unsigned long long* st = a.trace + (size_t)2 * (5 * a.L + 1) * a.nsteps;
    st[0] = (unsigned long long)t_full;
    st[1] = (unsigned long long)t_x;
    st[2] = (unsigned long long)t_slots;
    st[3] = (unsigned long long)t_bar;
    st[4] = (unsigned long long)(clock64() - t_begin);
    st[5] = seq;
    st[6] = (unsigned long long)t_flush;
  }
  if (blockIdx.x == 0 && tid < B) {
    a.next_token[s_slot[tid]] = s_tok[tid];
    a.seq_len[s_slot[tid]] = s_pos0[tid] + a.nsteps;
  }
}

size_t ringb_smem_bytes(int G, int D) {
  const size_t att = (size_t)G * D + (size_t)RW * G * (D + 2) + (size_t)RW * G * (TP + 1);
  return (size_t)NSLOT * MSLOT_BYTES + att * sizeof(float);
}
// static shared memory of the kernel (s_acc, s_rope and the small arrays), upper bound
size_t ringb_static_bytes(int D) { return (size_t)RW * NGB * 4 * 8 * 4 + (size_t)4 * 2 * (D / 2) * 4 + 1024; }

template <int D, int G, typename KVT, int NP, bool ONE = false>
cudaError_t launch_t(const MegaArgs& a, int grid, cudaStream_t s) {
  auto kern = decode_ring_batch_kernel<D, G, KVT, NP, ONE>;
  const size_t smem = ringb_smem_bytes(G, D);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  MegaArgs copy = a;
  void* params[] = {&copy};
  return cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3(RTHREADS), params, smem, s);
}

template <int D, int G, typename KVT>
cudaError_t launch_np(const MegaArgs& a, int grid, cudaStream_t s) {
  // batch 1: the two-part kernel with the second sequence absent (16 mantissa bits of the activations, like every
  // batched path) measured 0.630 ms per step, the three-part variant (24 bits) 0.650 -- LLMC_RINGB_ONE=1 selects it
  static const bool one = getenv("LLMC_RINGB_ONE") != nullptr && atoi(getenv("LLMC_RINGB_ONE")) != 0;
  if (a.B == 1 && one) return launch_t<D, G, KVT, 1, true>(a, grid, s);
  static const bool force2 = getenv("LLMC_RINGB_NP2") != nullptr;   // experiment knob: the two-pair kernel for batch 2 as well
  return (a.B <= 2 && !force2) ? launch_t<D, G, KVT, 1>(a, grid, s) : launch_t<D, G, KVT, 2>(a, grid, s);
}

template <typename KVT>
cudaError_t launch_kvt(const MegaArgs& a, int grid, cudaStream_t s) {
  const int G = a.q_heads / a.kv_heads;
  // the shapes of the checkpoints this path is built for: head size 64 or 128, 4 query heads per kv head
  if (a.head_dim == 64 && G == 4) return launch_np<64, 4, KVT>(a, grid, s);
  if (a.head_dim == 128 && G == 4) return launch_np<128, 4, KVT>(a, grid, s);
  return cudaErrorInvalidValue;
}

bool seg_ok(int K) { return K == SEG || K == 2 * SEG || K == 4 * SEG; }

}  // namespace

bool decode_ring_batch_supported(int B, int H, int QD, int KD, int I, int V, int q_heads, int kv_heads, int head_dim,
                                 int sm_count) {
  if (B < 1 || B > 4) return false;
  if (kv_heads < 1 || q_heads % kv_heads || q_heads / kv_heads != 4) return false;
  if (!(head_dim == 64 || head_dim == 128)) return false;
  if (H != SEG || !seg_ok(QD) || !seg_ok(I)) return false;
  if (((QD + 2 * KD) / 2) % 4 || V % 8 || I % 64) return false;
  if (H / 8 > 2 * sm_count || B * kv_heads > sm_count || sm_count > 192) return false;
  return ringb_smem_bytes(4, head_dim) + ringb_static_bytes(head_dim) <= 227 * 1024;
}

// scratch floats for the split-attention records of 4 sequences
size_t decode_ring_batch_scratch_floats(int q_heads, int head_dim) {
  return (size_t)4 * q_heads * RING_MAX_SPLIT * (head_dim + 2);
}

cudaError_t launch_decode_ring_batch(const MegaArgs& a, int sm_count, cudaStream_t s) {
  if (!decode_ring_batch_supported(a.B, a.H, a.QD, a.KD, a.I, a.V, a.q_heads, a.kv_heads, a.head_dim, sm_count))
    return cudaErrorInvalidValue;
  return a.kv_fp32 ? launch_kvt<float>(a, sm_count, s) : launch_kvt<uint16_t>(a, sm_count, s);
}

}  // namespace llmc
