// engine_internal.h -- the engine object behind the opaque llmc_engine handle.
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/llmc.h"
#include "common.cuh"
#include "kernels.h"

namespace llmc {
constexpr int MAX_STEPS_PER_CALL = 4096;
struct LayerW {
  uint16_t *g1, *wqkv, *wo, *g2, *wgu, *wd;
};
}  // namespace llmc

struct llmc_engine;
namespace llmc {
// checkpoint.cu: fill the (already laid out) weight blob from a safetensors file or directory of shards
int load_checkpoint(llmc_engine* e, const char* path, std::string* err);
// RoPE frequency scaling (Hugging Face rope_type "llama3"); kind 0 = none
struct RopeScaling {
  int kind = 0;
  float factor = 1.f, low_freq_factor = 1.f, high_freq_factor = 4.f;
  int original_ctx = 8192;
};
// checkpoint.cu: rope_theta / rope_scaling from the config.json beside a checkpoint (absent file: nothing changes)
int read_checkpoint_config(const char* weights_path, RopeScaling* rs, float* rope_theta, std::string* err);
}  // namespace llmc

struct llmc_engine {
  int device = 0;
  llmc::ModelDims d{};
  llmc::RopeScaling rope{};
  uint64_t seed = 0;
  int max_slots = 8, max_ctx = 8192, page_size = 64, prefill_chunk = 2048, prefill_impl = 0;
  bool use_graph = true, use_pdl = true;
  int pages_per_seq = 0, num_pages = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaDeviceProp prop{};
  // weights
  uint16_t* wblob = nullptr;
  size_t wblob_elems = 0;
  std::vector<llmc::LayerW> L;
  uint16_t *emb = nullptr, *lmhead = nullptr, *gf = nullptr;
  // kv
  unsigned char *kpool = nullptr, *vpool = nullptr;  // bf16 or fp32 elements
  size_t kv_layer_stride = 0;                        // elements per layer
  int kv_fp32 = 1, kv_elem = 4;
  bool split_prefill = true;
  int split_mask = 15;  // LLMC_PREFILL_SPLIT_MASK: which prefill GEMMs take the hi+lo pair (1 QKV, 2 O, 4 gate/up, 8 down)
  int* d_block_table = nullptr;
  std::vector<int> h_block_table;
  std::vector<int> free_pages;
  std::vector<int> slot_npages;
  std::vector<char> slot_used;
  std::vector<int> page_ref;   // sequences holding each KV page (llmc_seq_fork shares full pages read-only)
  std::vector<int> h_seq_len;
  int *d_seq_len = nullptr, *d_next_token = nullptr;
  float *rope_cos = nullptr, *rope_sin = nullptr;
  // decode workspace
  float *x = nullptr, *q = nullptr, *attn = nullptr, *act = nullptr, *logits = nullptr;
  int* d_slots = nullptr;
  float* part_val = nullptr;
  int* part_idx = nullptr;
  unsigned *lm_counter = nullptr, *attn_counters = nullptr;
  float* attn_scratch = nullptr;
  int *d_out_tokens = nullptr, *d_step_counter = nullptr, *d_dummy_counter = nullptr, *d_pslot = nullptr;
  // prefill workspace
  int* p_tokens = nullptr;
  float *px = nullptr, *pqkv = nullptr, *pq = nullptr;
  uint16_t *pxn = nullptr, *pattn = nullptr, *pact = nullptr;
  uint16_t *pxn_lo = nullptr, *pattn_lo = nullptr, *pact_lo = nullptr;  // split-bf16 low halves (or NULL)
  uint16_t* pattn_ws = nullptr;  // split K/V workspace of the tensor-core prefill attention (one layer, max_ctx positions)
  // one-kernel batch-1 decode
  bool use_mega = true;
  int wide_min_batch = 2;  // batches >= this use the wide (GEMM) decode path (LLMC_WIDE_MIN_BATCH=9: GEMV kernels up to 8)
  bool use_ring = false;  // persistent batch-1 decode through the TMA shared-memory ring (decode_ring.cu)
  bool use_ring_mma = false;  // ... with the slot arithmetic on mma.sync (decode_ring_mma.cu)
  bool use_ringb = false;  // 2..4 sequences: one persistent launch per decode call (decode_ring_batch.cu)
  float* ringb_scratch = nullptr;   // split-attention records of 4 sequences
  float* ringb_part_val = nullptr;  // [4][SMs] argmax candidates
  int* ringb_part_idx = nullptr;
  unsigned* ringb_counters = nullptr;   // [4][kv_heads], zero between launches
  uint16_t *ringb_attn_hi = nullptr, *ringb_attn_lo = nullptr, *ringb_act_hi = nullptr, *ringb_act_lo = nullptr, *ringb_attn_lo2 = nullptr, *ringb_act_lo2 = nullptr, *ringb_xn_hi = nullptr, *ringb_xn_lo = nullptr;
  unsigned long long* ringb_ss_acc = nullptr;   // [4 slots][4 sequences]   // [4][QD], [4][I] bf16 planes
  bool use_flow = false;  // ... with tagged-word hand-overs instead of grid barriers (decode_flow.cu)
  unsigned long long* flow_words = nullptr;   // the tagged-word buffers of decode_flow.cu
  size_t flow_nwords = 0;
  unsigned flow_tag = 1;                       // next unused hand-over tag
  llmc::MegaLayer* d_mega_layers = nullptr;
  float* mega_scratch = nullptr;
  int mega_max_tiles = 0;
  float* mega_part_val = nullptr;
  int* mega_part_idx = nullptr;
  unsigned *mega_barrier = nullptr, *mega_counters = nullptr;
  unsigned long long* mega_trace = nullptr;  // LLMC_MEGA_TRACE=1: per-barrier times of the last launch
  std::map<int, cudaGraphExec_t> graphs;  // key = (B * 2 + want_logits) * 16 + nsplit
  std::mutex mu;
  std::string last_error;
  llmc_stats stats{};
  std::vector<void*> allocs;
};

