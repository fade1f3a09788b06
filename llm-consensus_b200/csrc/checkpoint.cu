// checkpoint.cu -- load a Llama-family safetensors checkpoint (one file, or a directory of shards) into the
// engine's HBM layout (behind llmc_config.weights_path / the CLI's --weights).
//
// File format (safetensors): u64 little-endian header length, a JSON object {name: {"dtype", "shape",
// "data_offsets": [begin, end]}, ...} (+ "__metadata__"), then the raw tensor bytes.  Tensor names follow the
// Hugging Face Llama convention (model.embed_tokens.weight, model.layers.N.self_attn.q_proj.weight, ...,
// model.norm.weight, lm_head.weight); q/k are in the rotate-half RoPE layout the kernels use, so no permutation is
// needed.  BF16 tensors are copied as they are; F32 / F16 tensors are rounded to bf16 (nearest-even) on the host.
// The only re-arrangement is the engine's own packing: q, k, v rows concatenated into one matrix, gate / up rows
// interleaved in blocks of 64 (common.cuh), done with strided device copies.
//
// No counterpart in the reference (/root/reference/internal/provider/provider.go:13-21 is the boundary; the
// reference's backends hold no weights).
#include <dirent.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "engine_internal.h"

namespace llmc {

namespace {

struct TensorRef {
  std::string dtype;
  std::vector<int64_t> shape;
  const unsigned char* data = nullptr;
  size_t bytes = 0;
};

struct Shard {
  int fd = -1;
  unsigned char* map = nullptr;
  size_t size = 0;
  ~Shard() {
    if (map) munmap(map, size);
    if (fd >= 0) close(fd);
  }
};

// ---- a JSON reader for the header: objects, arrays, strings, numbers, true/false/null
struct Json {
  const char* p;
  const char* end;
  std::string err;
  void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
  bool lit(char c) {
    ws();
    if (p < end && *p == c) { ++p; return true; }
    return false;
  }
  bool str(std::string* out) {
    ws();
    if (p >= end || *p != '"') return fail("expected a string");
    ++p;
    out->clear();
    while (p < end && *p != '"') {
      if (*p == '\\' && p + 1 < end) {
        ++p;
        switch (*p) {
          case 'n': out->push_back('\n'); break;
          case 't': out->push_back('\t'); break;
          case 'u': if (end - p > 4) { out->push_back('?'); p += 4; } break;   // names are ASCII; value not needed
          default: out->push_back(*p);
        }
        ++p;
      } else {
        out->push_back(*p++);
      }
    }
    if (p >= end) return fail("unterminated string");
    ++p;
    return true;
  }
  bool num(int64_t* out) {
    ws();
    char* e = nullptr;
    const long long v = strtoll(p, &e, 10);
    if (e == p) return fail("expected a number");
    p = e;
    *out = v;
    return true;
  }
  bool skip() {   // any value
    ws();
    if (p >= end) return fail("truncated");
    if (*p == '"') { std::string s; return str(&s); }
    if (*p == '{' || *p == '[') {
      const char open = *p, close = open == '{' ? '}' : ']';
      ++p;
      if (lit(close)) return true;
      do {
        if (open == '{') {
          std::string k;
          if (!str(&k) || !lit(':')) return fail("bad object");
        }
        if (!skip()) return false;
      } while (lit(','));
      return lit(close) || fail("unterminated container");
    }
    while (p < end && *p != ',' && *p != '}' && *p != ']') ++p;   // number / true / false / null
    return true;
  }
  bool fail(const char* m) { if (err.empty()) err = m; return false; }
};

bool parse_header(const unsigned char* base, size_t file_size, std::map<std::string, TensorRef>* out, std::string* err) {
  if (file_size < 8) { *err = "file shorter than the 8-byte header length"; return false; }
  uint64_t n = 0;
  memcpy(&n, base, 8);
  if (n > file_size - 8) { *err = "header length exceeds the file"; return false; }
  const unsigned char* data0 = base + 8 + n;
  const size_t data_size = file_size - 8 - (size_t)n;
  Json j{reinterpret_cast<const char*>(base + 8), reinterpret_cast<const char*>(base + 8 + n), ""};
  if (!j.lit('{')) { *err = "header is not a JSON object"; return false; }
  if (j.lit('}')) return true;
  do {
    std::string name;
    if (!j.str(&name) || !j.lit(':')) { *err = "bad header: " + j.err; return false; }
    if (name == "__metadata__") {
      if (!j.skip()) { *err = "bad header metadata: " + j.err; return false; }
      continue;
    }
    TensorRef t;
    int64_t b = -1, e = -1;
    if (!j.lit('{')) { *err = "bad entry for " + name; return false; }
    do {
      std::string key;
      if (!j.str(&key) || !j.lit(':')) { *err = "bad entry for " + name; return false; }
      if (key == "dtype") {
        if (!j.str(&t.dtype)) { *err = "bad dtype for " + name; return false; }
      } else if (key == "shape") {
        if (!j.lit('[')) { *err = "bad shape for " + name; return false; }
        if (!j.lit(']')) {
          do {
            int64_t v;
            if (!j.num(&v)) { *err = "bad shape for " + name; return false; }
            t.shape.push_back(v);
          } while (j.lit(','));
          if (!j.lit(']')) { *err = "bad shape for " + name; return false; }
        }
      } else if (key == "data_offsets") {
        if (!j.lit('[') || !j.num(&b) || !j.lit(',') || !j.num(&e) || !j.lit(']')) { *err = "bad data_offsets for " + name; return false; }
      } else if (!j.skip()) {
        *err = "bad entry for " + name;
        return false;
      }
    } while (j.lit(','));
    if (!j.lit('}')) { *err = "bad entry for " + name; return false; }
    if (b < 0 || e < b || (uint64_t)e > data_size) { *err = "data_offsets of " + name + " outside the file"; return false; }
    t.data = data0 + b;
    t.bytes = (size_t)(e - b);
    (*out)[name] = t;
  } while (j.lit(','));
  if (!j.lit('}')) { *err = "unterminated header"; return false; }
  return true;
}

bool open_shard(const std::string& path, Shard* sh, std::string* err) {
  sh->fd = open(path.c_str(), O_RDONLY);
  if (sh->fd < 0) { *err = "open " + path + ": " + strerror(errno); return false; }
  struct stat st;
  if (fstat(sh->fd, &st) != 0) { *err = "stat " + path + ": " + strerror(errno); return false; }
  sh->size = (size_t)st.st_size;
  void* m = mmap(nullptr, sh->size, PROT_READ, MAP_PRIVATE, sh->fd, 0);
  if (m == MAP_FAILED) { *err = "mmap " + path + ": " + strerror(errno); return false; }
  sh->map = reinterpret_cast<unsigned char*>(m);
  return true;
}

uint16_t f32_to_bf16_host(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7f800000u) == 0x7f800000u && (u & 0x007fffffu)) return (uint16_t)((u >> 16) | 0x0040u);   // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
float f16_to_f32_host(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
  uint32_t u;
  if (exp == 0) {
    if (man == 0) u = sign;
    else {
      int e = -1;
      uint32_t m = man;
      do { ++e; m <<= 1; } while (!(m & 0x400u));
      u = sign | ((uint32_t)(127 - 15 - e) << 23) | ((m & 0x3ffu) << 13);
    }
  } else if (exp == 31) {
    u = sign | 0x7f800000u | (man << 13);
  } else {
    u = sign | ((exp + 112u) << 23) | (man << 13);
  }
  float f;
  memcpy(&f, &u, 4);
  return f;
}

struct Loader {
  std::vector<std::unique_ptr<Shard>> shards;
  std::map<std::string, TensorRef> tensors;
  std::vector<uint16_t> tmp;   // converted rows
  std::string err;

  bool open_path(const std::string& path) {
    struct stat st;
    if (stat(path.c_str(), &st) != 0) { err = "weights_path " + path + ": " + strerror(errno); return false; }
    std::vector<std::string> files;
    if (S_ISDIR(st.st_mode)) {
      DIR* d = opendir(path.c_str());
      if (!d) { err = "opendir " + path + ": " + strerror(errno); return false; }
      while (struct dirent* de = readdir(d)) {
        const std::string n = de->d_name;
        if (n.size() > 12 && n.compare(n.size() - 12, 12, ".safetensors") == 0) files.push_back(path + "/" + n);
      }
      closedir(d);
      std::sort(files.begin(), files.end());
      if (files.empty()) { err = "no .safetensors file in " + path; return false; }
    } else {
      files.push_back(path);
    }
    for (const std::string& f : files) {
      shards.emplace_back(new Shard());
      if (!open_shard(f, shards.back().get(), &err)) return false;
      std::string e2;
      if (!parse_header(shards.back()->map, shards.back()->size, &tensors, &e2)) { err = f + ": " + e2; return false; }
    }
    return true;
  }

  // the tensor as bf16 rows (pointer into the mapping, or into `tmp` after conversion)
  const uint16_t* get(const std::string& name, int64_t rows, int64_t cols) {
    auto it = tensors.find(name);
    if (it == tensors.end()) { err = "checkpoint has no tensor " + name; return nullptr; }
    const TensorRef& t = it->second;
    const bool shape_ok = (cols == 0 && t.shape.size() == 1 && t.shape[0] == rows) ||
                          (cols > 0 && t.shape.size() == 2 && t.shape[0] == rows && t.shape[1] == cols);
    if (!shape_ok) {
      std::string got;
      for (int64_t v : t.shape) got += (got.empty() ? "" : "x") + std::to_string(v);
      err = name + ": shape " + got + ", expected " + std::to_string(rows) + (cols ? "x" + std::to_string(cols) : "");
      return nullptr;
    }
    const size_t n = (size_t)rows * (size_t)(cols ? cols : 1);
    if (t.dtype == "BF16") {
      if (t.bytes != n * 2) { err = name + ": byte size does not match its shape"; return nullptr; }
      return reinterpret_cast<const uint16_t*>(t.data);
    }
    tmp.resize(n);
    if (t.dtype == "F32") {
      if (t.bytes != n * 4) { err = name + ": byte size does not match its shape"; return nullptr; }
      for (size_t i = 0; i < n; ++i) {
        float f;
        memcpy(&f, t.data + 4 * i, 4);
        tmp[i] = f32_to_bf16_host(f);
      }
    } else if (t.dtype == "F16") {
      if (t.bytes != n * 2) { err = name + ": byte size does not match its shape"; return nullptr; }
      for (size_t i = 0; i < n; ++i) {
        uint16_t h;
        memcpy(&h, t.data + 2 * i, 2);
        tmp[i] = f32_to_bf16_host(f16_to_f32_host(h));
      }
    } else {
      err = name + ": unsupported dtype " + t.dtype + " (BF16, F16, F32)";
      return nullptr;
    }
    return tmp.data();
  }
};

#define CK_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      *err = std::string(#expr) + ": " + cudaGetErrorString(_e);                        \
      return LLMC_ERR_CUDA;                                                             \
    }                                                                                   \
  } while (0)

}  // namespace

// ---- config.json beside the checkpoint: only rope_theta and rope_scaling are read.  Real Llama-3.1 / 3.2
// checkpoints declare rope_scaling {"rope_type": "llama3", "factor", "low_freq_factor", "high_freq_factor",
// "original_max_position_embeddings"}, which rescales inv_freq at every position; ignoring it would load without
// error and then disagree with transformers from the first token.
namespace {
// position just after `"key"` + ':' at nesting depth 1 of the object starting at s[from] == '{', or npos
size_t find_key(const std::string& s, size_t from, const std::string& key) {
  int depth = 0;
  bool in_str = false;
  for (size_t i = from; i < s.size(); ++i) {
    const char c = s[i];
    if (in_str) {
      if (c == '\\') ++i;
      else if (c == '"') in_str = false;
      continue;
    }
    if (c == '"') {
      if (depth == 1 && s.compare(i + 1, key.size(), key) == 0 && s[i + 1 + key.size()] == '"') {
        size_t j = i + 2 + key.size();
        while (j < s.size() && isspace((unsigned char)s[j])) ++j;
        if (j < s.size() && s[j] == ':') {
          ++j;
          while (j < s.size() && isspace((unsigned char)s[j])) ++j;
          return j;
        }
      }
      in_str = true;
    } else if (c == '{' || c == '[') {
      ++depth;
    } else if (c == '}' || c == ']') {
      if (--depth == 0) break;
    }
  }
  return std::string::npos;
}
bool number_at(const std::string& s, size_t at, double* v) {
  if (at == std::string::npos) return false;
  char* end = nullptr;
  *v = strtod(s.c_str() + at, &end);
  return end != s.c_str() + at;
}
}  // namespace

int read_checkpoint_config(const char* weights_path, RopeScaling* rs, float* rope_theta, std::string* err) {
  std::string dir = weights_path;
  struct stat st;
  if (stat(dir.c_str(), &st) != 0) return LLMC_OK;   // the loader reports a missing path
  if (!S_ISDIR(st.st_mode)) {
    const size_t sl = dir.rfind('/');
    dir = sl == std::string::npos ? "." : dir.substr(0, sl);
  }
  const std::string path = dir + "/config.json";
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return LLMC_OK;   // a bare .safetensors file: the caller's llmc_config is all there is
  std::string js;
  char buf[4096];
  size_t n;
  while ((n = fread(buf, 1, sizeof buf, f)) > 0) js.append(buf, n);
  fclose(f);
  const size_t root = js.find('{');
  if (root == std::string::npos) { *err = path + ": not a JSON object"; return LLMC_ERR_INVALID; }
  double v;
  if (number_at(js, find_key(js, root, "rope_theta"), &v)) *rope_theta = (float)v;
  const size_t at = find_key(js, root, "rope_scaling");
  if (at == std::string::npos || js.compare(at, 4, "null") == 0) return LLMC_OK;
  if (js[at] != '{') { *err = path + ": rope_scaling is neither null nor an object"; return LLMC_ERR_INVALID; }
  size_t ty = find_key(js, at, "rope_type");
  if (ty == std::string::npos) ty = find_key(js, at, "type");
  std::string type = "default";
  if (ty != std::string::npos && js[ty] == '"') type = js.substr(ty + 1, js.find('"', ty + 1) - ty - 1);
  if (type == "default") return LLMC_OK;
  if (type != "llama3") {
    *err = path + ": rope_scaling type \"" + type + "\" is not supported (default, llama3); the checkpoint would load "
           "but compute different logits than its reference implementation";
    return LLMC_ERR_INVALID;
  }
  double fa, lo, hi, oc;
  if (!number_at(js, find_key(js, at, "factor"), &fa) || !number_at(js, find_key(js, at, "low_freq_factor"), &lo) ||
      !number_at(js, find_key(js, at, "high_freq_factor"), &hi) ||
      !number_at(js, find_key(js, at, "original_max_position_embeddings"), &oc) || hi == lo) {
    *err = path + ": rope_scaling of type llama3 needs factor, low_freq_factor, high_freq_factor (distinct) and "
           "original_max_position_embeddings";
    return LLMC_ERR_INVALID;
  }
  rs->kind = LLMC_ROPE_LLAMA3;
  rs->factor = (float)fa; rs->low_freq_factor = (float)lo; rs->high_freq_factor = (float)hi; rs->original_ctx = (int)oc;
  return LLMC_OK;
}

// Fills the weight pointers of an engine whose blob is already allocated and laid out (engine.cu: init_weights).
int load_checkpoint(llmc_engine* e, const char* path, std::string* err) {
  const ModelDims& d = e->d;
  const int H = d.hidden, QD = d.qd(), KD = d.kd(), I = d.inter, V = d.vocab;
  Loader ld;
  if (!ld.open_path(path)) { *err = ld.err; return LLMC_ERR_INVALID; }
  auto put = [&](uint16_t* dst, const std::string& name, int64_t rows, int64_t cols) -> int {
    const uint16_t* src = ld.get(name, rows, cols);
    if (!src) { *err = ld.err; return LLMC_ERR_INVALID; }
    CK_CUDA(cudaMemcpy(dst, src, (size_t)rows * (size_t)(cols ? cols : 1) * 2, cudaMemcpyHostToDevice));
    return LLMC_OK;
  };
  // logical row j of gate (half 0) / up (half 1) -> packed row (j / 64) * 128 + half * 64 + j % 64
  auto put_gate_up = [&](uint16_t* dst, const std::string& name, int half) -> int {
    const uint16_t* src = ld.get(name, I, H);
    if (!src) { *err = ld.err; return LLMC_ERR_INVALID; }
    const size_t blk = (size_t)GU_BLK * H * 2;
    CK_CUDA(cudaMemcpy2D(reinterpret_cast<unsigned char*>(dst) + (size_t)half * blk, 2 * blk, src, blk, blk,
                         (size_t)I / GU_BLK, cudaMemcpyHostToDevice));
    return LLMC_OK;
  };
  for (int l = 0; l < d.layers; ++l) {
    const LayerW& w = e->L[l];
    const std::string p = "model.layers." + std::to_string(l) + ".";
    if (int rc = put(w.g1, p + "input_layernorm.weight", H, 0)) return rc;
    if (int rc = put(w.wqkv, p + "self_attn.q_proj.weight", QD, H)) return rc;
    if (int rc = put(w.wqkv + (size_t)QD * H, p + "self_attn.k_proj.weight", KD, H)) return rc;
    if (int rc = put(w.wqkv + (size_t)(QD + KD) * H, p + "self_attn.v_proj.weight", KD, H)) return rc;
    if (int rc = put(w.wo, p + "self_attn.o_proj.weight", H, QD)) return rc;
    if (int rc = put(w.g2, p + "post_attention_layernorm.weight", H, 0)) return rc;
    if (int rc = put_gate_up(w.wgu, p + "mlp.gate_proj.weight", 0)) return rc;
    if (int rc = put_gate_up(w.wgu, p + "mlp.up_proj.weight", 1)) return rc;
    if (int rc = put(w.wd, p + "mlp.down_proj.weight", H, I)) return rc;
  }
  if (int rc = put(e->emb, "model.embed_tokens.weight", V, H)) return rc;
  if (!d.tied)
    if (int rc = put(e->lmhead, "lm_head.weight", V, H)) return rc;
  if (int rc = put(e->gf, "model.norm.weight", H, 0)) return rc;
  return LLMC_OK;
}

}  // namespace llmc
