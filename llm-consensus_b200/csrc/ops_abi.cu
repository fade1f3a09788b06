// ops_abi.cu -- kernel-level test/bench hooks (include/llmc_ops.h): run one kernel of the hot
// path on host buffers.  Test plumbing around the product kernels; no arithmetic of its own.
#include <math.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/llmc_ops.h"
#include "engine_internal.h"

using namespace llmc;

namespace {

thread_local std::string g_ops_error;

int ofail(int code, const std::string& m) {
  g_ops_error = m;
  return code;
}

#define O_TRY(expr)                                                                                              \
  do {                                                                                                           \
    cudaError_t _e = (expr);                                                                                     \
    if (_e != cudaSuccess)                                                                                       \
      return ofail(_e == cudaErrorMemoryAllocation ? LLMC_ERR_OOM : LLMC_ERR_CUDA,                               \
                   std::string(#expr) + ": " + cudaGetErrorString(_e) + " (ops_abi.cu:" + std::to_string(__LINE__) + ")"); \
  } while (0)

// device buffers freed on scope exit
struct Pool {
  std::vector<void*> p;
  ~Pool() {
    for (void* q : p) cudaFree(q);
  }
  template <typename T>
  cudaError_t get(T** out, size_t n, const void* host = nullptr) {
    void* v = nullptr;
    cudaError_t e = cudaMalloc(&v, n * sizeof(T) + 256);
    if (e != cudaSuccess) return e;
    p.push_back(v);
    *out = reinterpret_cast<T*>(v);
    if (host) return cudaMemcpy(v, host, n * sizeof(T), cudaMemcpyHostToDevice);
    return cudaMemset(v, 0, n * sizeof(T));
  }
};

int batch_tile(int B) { return B <= 1 ? 1 : (B <= 2 ? 2 : (B <= 4 ? 4 : 8)); }

// scrambled page assignment so that a kernel that ignores the block table fails the test
void scrambled_table(int B, int pages_per_seq, std::vector<int>& bt) {
  const int n = B * pages_per_seq;
  bt.resize(n);
  for (int i = 0; i < n; ++i) bt[i] = (int)(((uint64_t)i * 7919u + 13u) % (uint64_t)n);
  // 7919 is prime; if it divides n fall back to a reversal
  std::vector<char> seen(n, 0);
  bool ok = true;
  for (int i = 0; i < n; ++i) {
    if (seen[bt[i]]) ok = false;
    seen[bt[i]] = 1;
  }
  if (!ok)
    for (int i = 0; i < n; ++i) bt[i] = n - 1 - i;
}

}  // namespace

extern "C" {

const char* llmc_ops_last_error(void) { return g_ops_error.c_str(); }

int llmc_op_gemv(int device, int N, int K, int B, const uint16_t* W, const float* x, int norm,
                 const uint16_t* gain, float eps, int resid, float* y) {
  if (!W || !x || !y || B < 1 || B > MAX_GEMV_BATCH || (norm && !gain)) return ofail(LLMC_ERR_INVALID, "llmc_op_gemv: bad argument");
  O_TRY(cudaSetDevice(device));
  Pool pool;
  uint16_t *dW, *dg = nullptr;
  float *dx, *dy;
  O_TRY(pool.get(&dW, (size_t)N * K, W));
  O_TRY(pool.get(&dx, (size_t)B * K, x));
  O_TRY(pool.get(&dy, (size_t)B * N, y));
  if (norm) O_TRY(pool.get(&dg, (size_t)K, gain));
  GemvArgs a;
  memset(&a, 0, sizeof(a));
  a.W = dW; a.N = N; a.K = K; a.B = B;
  a.x = dx; a.x_stride = K; a.gain = dg; a.eps = eps;
  a.y = dy; a.y_stride = N;
  O_TRY(launch_gemv_any(resid ? EPI_RESID : EPI_STORE, norm ? PRO_NORM : PRO_PLAIN, a, false, 0));
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaMemcpy(y, dy, (size_t)B * N * sizeof(float), cudaMemcpyDeviceToHost));
  return LLMC_OK;
}

int llmc_op_gemv_swiglu(int device, int I, int K, int B, const uint16_t* Wg, const uint16_t* Wu, const float* x,
                        const uint16_t* gain, float eps, float* act) {
  if (!Wg || !Wu || !x || !gain || !act || B < 1 || B > MAX_GEMV_BATCH || I % GU_BLK) return ofail(LLMC_ERR_INVALID, "llmc_op_gemv_swiglu: bad argument");
  O_TRY(cudaSetDevice(device));
  Pool pool;
  std::vector<uint16_t> packed((size_t)2 * I * K);
  for (int j = 0; j < I; ++j) {
    memcpy(&packed[(size_t)gu_row_gate(j) * K], Wg + (size_t)j * K, (size_t)K * 2);
    memcpy(&packed[(size_t)gu_row_up(j) * K], Wu + (size_t)j * K, (size_t)K * 2);
  }
  uint16_t *dW, *dg;
  float *dx, *dy;
  O_TRY(pool.get(&dW, packed.size(), packed.data()));
  O_TRY(pool.get(&dx, (size_t)B * K, x));
  O_TRY(pool.get(&dy, (size_t)B * I));
  O_TRY(pool.get(&dg, (size_t)K, gain));
  GemvArgs a;
  memset(&a, 0, sizeof(a));
  a.W = dW; a.N = 2 * I; a.K = K; a.B = B;
  a.x = dx; a.x_stride = K; a.gain = dg; a.eps = eps;
  a.y = dy; a.y_stride = I;
  O_TRY(launch_gemv_any(EPI_SWIGLU, PRO_NORM, a, false, 0));
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaMemcpy(act, dy, (size_t)B * I * sizeof(float), cudaMemcpyDeviceToHost));
  return LLMC_OK;
}

int llmc_op_lmhead(int device, int V, int K, int B, const uint16_t* W, const float* x, const uint16_t* gain,
                   float eps, float* logits, int32_t* tok) {
  if (!W || !x || !gain || !tok || B < 1 || B > MAX_GEMV_BATCH) return ofail(LLMC_ERR_INVALID, "llmc_op_lmhead: bad argument");
  O_TRY(cudaSetDevice(device));
  Pool pool;
  uint16_t *dW, *dg;
  float *dx, *dl, *pv;
  int *pi, *slots, *next, *seq, *outtok, *step;
  unsigned* ctr;
  const int nb = gemv_lmhead_blocks(V) > gemv_mma_lmhead_blocks(V) ? gemv_lmhead_blocks(V) : gemv_mma_lmhead_blocks(V);
  O_TRY(pool.get(&dW, (size_t)V * K, W));
  O_TRY(pool.get(&dx, (size_t)B * K, x));
  O_TRY(pool.get(&dg, (size_t)K, gain));
  O_TRY(pool.get(&dl, (size_t)B * V));
  O_TRY(pool.get(&pv, (size_t)B * nb));
  O_TRY(pool.get(&pi, (size_t)B * nb));
  std::vector<int> hs(B);
  for (int b = 0; b < B; ++b) hs[b] = B - 1 - b;  // slot != batch row on purpose
  O_TRY(pool.get(&slots, (size_t)B, hs.data()));
  O_TRY(pool.get(&next, (size_t)B));
  O_TRY(pool.get(&seq, (size_t)B));
  O_TRY(pool.get(&outtok, (size_t)B * 4));
  O_TRY(pool.get(&step, 1));
  O_TRY(pool.get(&ctr, 1));
  GemvArgs a;
  memset(&a, 0, sizeof(a));
  a.W = dW; a.N = V; a.K = K; a.B = B;
  a.x = dx; a.x_stride = K; a.gain = dg; a.eps = eps;
  a.slots = slots;
  a.y = logits ? dl : nullptr; a.y_stride = V;
  a.part_val = pv; a.part_idx = pi; a.counter = ctr;
  a.next_token_out = next; a.seq_len_out = seq; a.seq_inc = 1;
  a.out_tokens = outtok; a.out_stride = 4; a.step_counter = step;
  // two launches: the second checks that counter/step state is left re-launchable
  for (int rep = 0; rep < 2; ++rep) O_TRY(launch_gemv_any(EPI_LMHEAD, PRO_NORM, a, false, 0));
  O_TRY(cudaDeviceSynchronize());
  std::vector<int> ho((size_t)B * 4), hn(B), hq(B);
  int hstep = 0;
  O_TRY(cudaMemcpy(ho.data(), outtok, ho.size() * sizeof(int), cudaMemcpyDeviceToHost));
  O_TRY(cudaMemcpy(hn.data(), next, (size_t)B * sizeof(int), cudaMemcpyDeviceToHost));
  O_TRY(cudaMemcpy(hq.data(), seq, (size_t)B * sizeof(int), cudaMemcpyDeviceToHost));
  O_TRY(cudaMemcpy(&hstep, step, sizeof(int), cudaMemcpyDeviceToHost));
  if (logits) O_TRY(cudaMemcpy(logits, dl, (size_t)B * V * sizeof(float), cudaMemcpyDeviceToHost));
  if (hstep != 2) return ofail(LLMC_ERR_STATE, "lm head: step counter is " + std::to_string(hstep) + ", expected 2");
  for (int b = 0; b < B; ++b) {
    if (ho[(size_t)b * 4] != ho[(size_t)b * 4 + 1] || hn[hs[b]] != ho[(size_t)b * 4] || hq[hs[b]] != 2)
      return ofail(LLMC_ERR_STATE, "lm head: inconsistent finalisation state for batch row " + std::to_string(b));
    tok[b] = ho[(size_t)b * 4];
  }
  return LLMC_OK;
}

int llmc_op_attn_decode(int device, int B, int q_heads, int kv_heads, int head_dim, int page_size, int nsplit,
                        int max_ctx, int kv_fp32, const int32_t* ctx, const float* q, const void* k, const void* v,
                        float* out) {
  if (!ctx || !q || !k || !v || !out || B < 1 || kv_heads < 1 || page_size < 32 || page_size % 32)
    return ofail(LLMC_ERR_INVALID, "llmc_op_attn_decode: bad argument");
  O_TRY(cudaSetDevice(device));
  const int D = head_dim, KD = kv_heads * D, QD = q_heads * D;
  const size_t es = kv_fp32 ? 4 : 2;
  const int pps = (max_ctx + page_size - 1) / page_size;
  std::vector<int> bt;
  scrambled_table(B, pps, bt);
  const size_t pool_elems = (size_t)B * pps * KD * page_size;
  // unwritten page slots hold NaN patterns: a kernel that reads past ctx poisons its output
  std::vector<unsigned char> hk(pool_elems * es, 0xFF), hv(pool_elems * es, 0xFF);
  const unsigned char* ks = reinterpret_cast<const unsigned char*>(k);
  const unsigned char* vs = reinterpret_cast<const unsigned char*>(v);
  int longest = 1;
  for (int b = 0; b < B; ++b) {
    if (ctx[b] < 1 || ctx[b] > max_ctx) return ofail(LLMC_ERR_INVALID, "llmc_op_attn_decode: ctx out of range");
    if (ctx[b] > longest) longest = ctx[b];
    for (int p = 0; p < ctx[b]; ++p) {
      const int page = bt[(size_t)b * pps + p / page_size], off = p % page_size;
      for (int h = 0; h < kv_heads; ++h) {
        const size_t dst = (((size_t)page * kv_heads + h) * page_size + off) * D;
        const size_t src = ((size_t)b * max_ctx + p) * KD + (size_t)h * D;
        memcpy(&hk[dst * es], ks + src * es, (size_t)D * es);
        memcpy(&hv[dst * es], vs + src * es, (size_t)D * es);
      }
    }
  }
  if (nsplit <= 0) nsplit = attn_decode_nsplit(longest);
  Pool pool;
  unsigned char *dk, *dv;
  float *dq, *dout, *scratch;
  int *dbt, *dslots, *dseq;
  unsigned* ctrs;
  std::vector<int> slots(B);
  for (int b = 0; b < B; ++b) slots[b] = b;
  O_TRY(pool.get(&dk, pool_elems * es, hk.data()));
  O_TRY(pool.get(&dv, pool_elems * es, hv.data()));
  O_TRY(pool.get(&dq, (size_t)B * QD, q));
  O_TRY(pool.get(&dout, (size_t)B * QD));
  O_TRY(pool.get(&scratch, attn_decode_scratch_floats(B, q_heads, nsplit, D)));
  O_TRY(pool.get(&dbt, bt.size(), bt.data()));
  O_TRY(pool.get(&dslots, (size_t)B, slots.data()));
  O_TRY(pool.get(&dseq, (size_t)B, ctx));
  O_TRY(pool.get(&ctrs, (size_t)B * kv_heads));
  AttnDecodeArgs a;
  memset(&a, 0, sizeof(a));
  a.q = dq; a.out = dout; a.kpool = dk; a.vpool = dv; a.kv_fp32 = kv_fp32;
  a.slots = dslots; a.seq_len = dseq; a.ctx_extra = 0;
  a.block_table = dbt; a.bt_stride = pps; a.page_size = page_size;
  a.q_heads = q_heads; a.kv_heads = kv_heads; a.head_dim = D; a.B = B; a.nsplit = nsplit;
  a.scratch = scratch; a.counters = ctrs;
  // twice: the counters must come back to zero for the second launch to be right
  for (int rep = 0; rep < 2; ++rep) {
    O_TRY(cudaMemset(dout, 0, (size_t)B * QD * sizeof(float)));
    O_TRY(launch_attn_decode(a, false, 0));
  }
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaMemcpy(out, dout, (size_t)B * QD * sizeof(float), cudaMemcpyDeviceToHost));
  return LLMC_OK;
}

int llmc_op_gemm(int device, int impl, int epi, int T, int N, int K, const uint16_t* A, const uint16_t* A_lo,
                 const uint16_t* W, void* out, void* out_lo) {
  if (!A || !W || !out || T < 1) return ofail(LLMC_ERR_INVALID, "llmc_op_gemm: bad argument");
  O_TRY(cudaSetDevice(device));
  Pool pool;
  std::vector<uint16_t> packed;
  const uint16_t* hw = W;
  if (epi == GEMM_SWIGLU_BF16) {
    const int I = N / 2;
    if (I % GU_BLK) return ofail(LLMC_ERR_INVALID, "llmc_op_gemm: I must be a multiple of 64");
    packed.resize((size_t)N * K);
    for (int j = 0; j < I; ++j) {
      memcpy(&packed[(size_t)gu_row_gate(j) * K], W + (size_t)j * K, (size_t)K * 2);
      memcpy(&packed[(size_t)gu_row_up(j) * K], W + (size_t)(I + j) * K, (size_t)K * 2);
    }
    hw = packed.data();
  }
  const size_t out_cols = epi == GEMM_SWIGLU_BF16 ? N / 2 : N;
  const size_t out_bytes = (size_t)T * out_cols * ((epi == GEMM_STORE_F32 || epi == GEMM_RESID_F32) ? 4 : 2);
  uint16_t *dA, *dAl = nullptr, *dW;
  unsigned char *dout, *dol = nullptr;
  O_TRY(pool.get(&dA, (size_t)T * K, A));
  if (A_lo) O_TRY(pool.get(&dAl, (size_t)T * K, A_lo));
  O_TRY(pool.get(&dW, (size_t)N * K, hw));
  O_TRY(pool.get(&dout, out_bytes, epi == GEMM_RESID_F32 ? out : nullptr));
  if (out_lo && epi == GEMM_SWIGLU_BF16) O_TRY(pool.get(&dol, out_bytes));
  if (impl == 3) O_TRY(launch_gemm_decode_tc(dA, dAl, dW, dout, dol, T, N, K, epi, 0));
  else if (impl == 0 || impl == 2) O_TRY(launch_gemm_tcgen05(dA, dAl, dW, dout, dol, T, N, K, epi, 0, impl == 2));
  else O_TRY(launch_gemm_simple(dA, dAl, dW, dout, dol, T, N, K, epi, 0));
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaMemcpy(out, dout, out_bytes, cudaMemcpyDeviceToHost));
  if (dol) O_TRY(cudaMemcpy(out_lo, dol, out_bytes, cudaMemcpyDeviceToHost));
  return LLMC_OK;
}

int llmc_op_gemm_bench(int device, int impl, int epi, int split, int T, int N, int K, int iters, float* ms_per_call) {
  if (T < 1 || iters < 1 || !ms_per_call) return ofail(LLMC_ERR_INVALID, "llmc_op_gemm_bench: bad argument");
  O_TRY(cudaSetDevice(device));
  Pool pool;
  uint16_t *dA, *dAl = nullptr, *dW;
  unsigned char *dout, *dol = nullptr;
  const size_t out_cols = epi == GEMM_SWIGLU_BF16 ? N / 2 : N;
  const size_t out_bytes = (size_t)T * out_cols * 4;
  O_TRY(pool.get(&dA, (size_t)T * K));
  if (split) O_TRY(pool.get(&dAl, (size_t)T * K));
  O_TRY(pool.get(&dW, (size_t)N * K));
  O_TRY(pool.get(&dout, out_bytes));
  if (split && epi == GEMM_SWIGLU_BF16) O_TRY(pool.get(&dol, out_bytes));
  // operands: a fixed non-trivial bit pattern (bf16 0x3c00.. small values); timing only
  O_TRY(cudaMemset(dA, 0x3c, (size_t)T * K * 2));
  O_TRY(cudaMemset(dW, 0x3c, (size_t)N * K * 2));
  cudaEvent_t e0, e1;
  O_TRY(cudaEventCreate(&e0));
  O_TRY(cudaEventCreate(&e1));
  auto run = [&]() {
    if (impl == 3) return launch_gemm_decode_tc(dA, dAl, dW, dout, dol, T, N, K, epi, 0);
    return (impl == 0 || impl == 2) ? launch_gemm_tcgen05(dA, dAl, dW, dout, dol, T, N, K, epi, 0, impl == 2)
                     : launch_gemm_simple(dA, dAl, dW, dout, dol, T, N, K, epi, 0);
  };
  for (int i = 0; i < 3; ++i) O_TRY(run());
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaEventRecord(e0, 0));
  for (int i = 0; i < iters; ++i) O_TRY(run());
  O_TRY(cudaEventRecord(e1, 0));
  O_TRY(cudaDeviceSynchronize());
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ms_per_call = ms / iters;
  return LLMC_OK;
}

int llmc_op_rmsnorm_rows(int device, int T, int H, const float* x, const uint16_t* gain, float eps, uint16_t* out,
                         uint16_t* out_lo) {
  if (!x || !gain || !out || T < 1 || H % 8) return ofail(LLMC_ERR_INVALID, "llmc_op_rmsnorm_rows: bad argument");
  O_TRY(cudaSetDevice(device));
  Pool pool;
  float* dx;
  uint16_t *dg, *dout, *dlo = nullptr;
  O_TRY(pool.get(&dx, (size_t)T * H, x));
  O_TRY(pool.get(&dg, (size_t)H, gain));
  O_TRY(pool.get(&dout, (size_t)T * H));
  if (out_lo) O_TRY(pool.get(&dlo, (size_t)T * H));
  O_TRY(launch_rmsnorm_rows(dx, dg, eps, T, H, dout, dlo, 0));
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaMemcpy(out, dout, (size_t)T * H * 2, cudaMemcpyDeviceToHost));
  if (out_lo) O_TRY(cudaMemcpy(out_lo, dlo, (size_t)T * H * 2, cudaMemcpyDeviceToHost));
  return LLMC_OK;
}

int llmc_op_prefill_attention(int device, int T, int p0, int q_heads, int kv_heads, int head_dim, int page_size,
                              float rope_theta, int kv_fp32, const float* qkv, const void* k_past, const void* v_past,
                              uint16_t* out, uint16_t* out_lo, float* q_out, void* k_new, void* v_new, int impl) {
  if (!qkv || !out || T < 1 || p0 < 0 || (p0 > 0 && (!k_past || !v_past)) || page_size < 32 || page_size % 32)
    return ofail(LLMC_ERR_INVALID, "llmc_op_prefill_attention: bad argument");
  O_TRY(cudaSetDevice(device));
  const int D = head_dim, KD = kv_heads * D, QD = q_heads * D, total = p0 + T;
  const size_t es = kv_fp32 ? 4 : 2;
  const int pps = (total + page_size - 1) / page_size;
  std::vector<int> bt;
  scrambled_table(1, pps, bt);
  const size_t pool_elems = (size_t)pps * KD * page_size;
  std::vector<unsigned char> hk(pool_elems * es, 0xFF), hv(pool_elems * es, 0xFF);
  const unsigned char* kps = reinterpret_cast<const unsigned char*>(k_past);
  const unsigned char* vps = reinterpret_cast<const unsigned char*>(v_past);
  for (int p = 0; p < p0; ++p) {
    const int page = bt[p / page_size], off = p % page_size;
    for (int h = 0; h < kv_heads; ++h) {
      const size_t dst = (((size_t)page * kv_heads + h) * page_size + off) * D;
      memcpy(&hk[dst * es], kps + ((size_t)p * KD + (size_t)h * D) * es, (size_t)D * es);
      memcpy(&hv[dst * es], vps + ((size_t)p * KD + (size_t)h * D) * es, (size_t)D * es);
    }
  }
  std::vector<float> hc((size_t)total * (D / 2)), hs((size_t)total * (D / 2));
  for (int p = 0; p < total; ++p)
    for (int i = 0; i < D / 2; ++i) {
      const double a = (double)p * pow((double)rope_theta, -2.0 * i / (double)D);
      hc[(size_t)p * (D / 2) + i] = (float)cos(a);
      hs[(size_t)p * (D / 2) + i] = (float)sin(a);
    }
  Pool pool;
  unsigned char *dk, *dv;
  uint16_t *dout, *dlo = nullptr;
  float *dqkv, *dq, *dc, *ds;
  int* dbt;
  O_TRY(pool.get(&dk, pool_elems * es, hk.data()));
  O_TRY(pool.get(&dv, pool_elems * es, hv.data()));
  O_TRY(pool.get(&dqkv, (size_t)T * (QD + 2 * KD), qkv));
  O_TRY(pool.get(&dq, (size_t)T * QD));
  O_TRY(pool.get(&dout, (size_t)T * QD));
  if (out_lo) O_TRY(pool.get(&dlo, (size_t)T * QD));
  O_TRY(pool.get(&dc, hc.size(), hc.data()));
  O_TRY(pool.get(&ds, hs.size(), hs.data()));
  O_TRY(pool.get(&dbt, bt.size(), bt.data()));
  O_TRY(launch_rope_append(dqkv, T, p0, 0, dc, ds, dbt, pps, dk, dv, kv_fp32, page_size, q_heads, kv_heads, D, dq, 0));
  uint16_t* dws = nullptr;
  if ((impl == 0 || impl == 2) && q_heads % kv_heads == 0 && prefill_attn_mma_supported(D, q_heads / kv_heads))
  {
    const size_t wsb = prefill_attn_mma_workspace_bytes(total, kv_heads, D);
    O_TRY(pool.get(&dws, wsb / 2));
    O_TRY(cudaMemset(dws, 0xFF, wsb));   // NaN poison: the kernel must never consume a byte it did not write
  }
  O_TRY(launch_prefill_attention(dq, T, p0, 0, dbt, pps, dk, dv, kv_fp32, page_size, q_heads, kv_heads, D, dout, dlo, dws,
                                 impl, 0));
  O_TRY(cudaDeviceSynchronize());
  O_TRY(cudaMemcpy(out, dout, (size_t)T * QD * 2, cudaMemcpyDeviceToHost));
  if (out_lo) O_TRY(cudaMemcpy(out_lo, dlo, (size_t)T * QD * 2, cudaMemcpyDeviceToHost));
  if (q_out) O_TRY(cudaMemcpy(q_out, dq, (size_t)T * QD * 4, cudaMemcpyDeviceToHost));
  if (k_new || v_new) {
    O_TRY(cudaMemcpy(hk.data(), dk, pool_elems * es, cudaMemcpyDeviceToHost));
    O_TRY(cudaMemcpy(hv.data(), dv, pool_elems * es, cudaMemcpyDeviceToHost));
    for (int t = 0; t < T; ++t) {
      const int p = p0 + t, page = bt[p / page_size], off = p % page_size;
      for (int h = 0; h < kv_heads; ++h) {
        const size_t src = (((size_t)page * kv_heads + h) * page_size + off) * D;
        const size_t dst = (size_t)t * KD + (size_t)h * D;
        if (k_new) memcpy(reinterpret_cast<unsigned char*>(k_new) + dst * es, &hk[src * es], (size_t)D * es);
        if (v_new) memcpy(reinterpret_cast<unsigned char*>(v_new) + dst * es, &hv[src * es], (size_t)D * es);
      }
    }
  }
  return LLMC_OK;
}

int64_t llmc_debug_tensor(llmc_engine* e, int layer, int kind, uint16_t* out, int64_t cap) {
  if (!e) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  const ModelDims& d = e->d;
  const int H = d.hidden, QD = d.qd(), KD = d.kd(), I = d.inter, V = d.vocab;
  const uint16_t* src = nullptr;
  int64_t rows = 0, cols = 0;
  int mode = 0;  // 0 contiguous, 1 gate rows, 2 up rows
  if (layer < 0) {
    if (kind == K_EMB) { src = e->emb; rows = V; cols = H; }
    else if (kind == K_LMHEAD) { src = e->lmhead; rows = V; cols = H; }
    else if (kind == K_FINAL_NORM) { src = e->gf; rows = 1; cols = H; }
  } else if (layer < d.layers) {
    const LayerW& w = e->L[layer];
    switch (kind) {
      case K_ATTN_NORM: src = w.g1; rows = 1; cols = H; break;
      case K_WQ: src = w.wqkv; rows = QD; cols = H; break;
      case K_WK: src = w.wqkv + (size_t)QD * H; rows = KD; cols = H; break;
      case K_WV: src = w.wqkv + (size_t)(QD + KD) * H; rows = KD; cols = H; break;
      case K_WO: src = w.wo; rows = H; cols = QD; break;
      case K_MLP_NORM: src = w.g2; rows = 1; cols = H; break;
      case K_WG: src = w.wgu; rows = I; cols = H; mode = 1; break;
      case K_WU: src = w.wgu; rows = I; cols = H; mode = 2; break;
      case K_WD: src = w.wd; rows = H; cols = I; break;
    }
  }
  if (!src) return LLMC_ERR_INVALID;
  const int64_t n = rows * cols;
  if (!out || cap < n) return n;
  if (cudaSetDevice(e->device) != cudaSuccess) return LLMC_ERR_CUDA;
  cudaStreamSynchronize(e->stream);
  if (mode == 0) {
    if (cudaMemcpy(out, src, (size_t)n * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return LLMC_ERR_CUDA;
  } else {
    for (int64_t j0 = 0; j0 < rows; j0 += GU_BLK) {
      const int r = mode == 1 ? gu_row_gate((int)j0) : gu_row_up((int)j0);
      if (cudaMemcpy(out + j0 * cols, src + (size_t)r * cols, (size_t)GU_BLK * cols * 2, cudaMemcpyDeviceToHost) !=
          cudaSuccess)
        return LLMC_ERR_CUDA;
    }
  }
  return n;
}

int llmc_debug_kv(llmc_engine* e, int slot, int layer, int pos, void* k, void* v) {
  if (!e || !k || !v) return LLMC_ERR_INVALID;
  std::lock_guard<std::mutex> lk(e->mu);
  const ModelDims& d = e->d;
  if (slot < 0 || slot >= e->max_slots || layer < 0 || layer >= d.layers || pos < 0 || pos >= e->h_seq_len[slot])
    return LLMC_ERR_INVALID;
  if (cudaSetDevice(e->device) != cudaSuccess) return LLMC_ERR_CUDA;
  cudaStreamSynchronize(e->stream);
  const int D = d.head_dim;
  const size_t es = e->kv_elem;
  const int page = e->h_block_table[(size_t)slot * e->pages_per_seq + pos / e->page_size], off = pos % e->page_size;
  for (int h = 0; h < d.kv_heads; ++h) {
    const size_t src = (size_t)layer * e->kv_layer_stride + (((size_t)page * d.kv_heads + h) * e->page_size + off) * D;
    if (cudaMemcpy(reinterpret_cast<unsigned char*>(k) + (size_t)h * D * es, e->kpool + src * es, (size_t)D * es,
                   cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(reinterpret_cast<unsigned char*>(v) + (size_t)h * D * es, e->vpool + src * es, (size_t)D * es,
                   cudaMemcpyDeviceToHost) != cudaSuccess)
      return LLMC_ERR_CUDA;
  }
  return LLMC_OK;
}

}  // extern "C"
