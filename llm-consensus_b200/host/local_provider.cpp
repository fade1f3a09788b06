// local_provider.cpp -- the local backend: implements Provider (provider.go:13-21) on top of the
// C ABI of include/llmc.h, shaped like the remote backends' QueryStream
// (/root/reference/internal/provider/openai.go:141-210): build the request, run it, call the
// callback synchronously on the caller's thread for every chunk, return
// Response{Model: req.Model, Content, Provider, Latency}.
#include <chrono>

#include "../../include/llmc.h"
#include "host.hpp"

namespace host {

namespace {

class Local : public Provider {
 public:
  Local(llmc_engine* e, const LocalConfig& cfg, int vocab) : e_(e), cfg_(cfg), vocab_(vocab) {}
  ~Local() override { llmc_engine_destroy(e_); }

  Error Query(const Context& ctx, const Request& req, Response* out) override {
    return QueryStream(ctx, req, nullptr, out);  // never called by the CLI path (runner.go:92, judge.go:96)
  }

  Error QueryStream(const Context& ctx, const Request& req, const StreamCallback& cb, Response* out) override {
    const auto start = std::chrono::steady_clock::now();
    std::vector<int32_t> ids = Tokenize(req.Prompt, vocab_);
    if (ids.empty()) ids.push_back('\n');
    if ((int)ids.size() + cfg_.max_new_tokens > cfg_.max_ctx)
      return Error::New("prompt of " + std::to_string(ids.size()) + " tokens plus " +
                        std::to_string(cfg_.max_new_tokens) + " new tokens exceeds the context of " +
                        std::to_string(cfg_.max_ctx));
    if (!ctx.Err().empty()) return Error::New(ctx.Err());
    int32_t slot = -1;
    if (llmc_seq_alloc(e_, &slot) != LLMC_OK) return fail("allocating a sequence slot");
    struct Free {
      llmc_engine* e;
      int32_t s;
      ~Free() { llmc_seq_free(e, s); }
    } guard{e_, slot};
    std::string content;
    auto emit = [&](const int32_t* t, size_t n) {
      const std::string chunk = Detokenize(t, n);
      content += chunk;
      if (cb) cb(chunk);
    };
    int32_t first = -1;
    if (llmc_prefill(e_, slot, ids.data(), (int32_t)ids.size(), &first, nullptr) != LLMC_OK) return fail("prefill");
    emit(&first, 1);
    std::vector<int32_t> toks(cfg_.chunk);
    for (int left = cfg_.max_new_tokens - 1; left > 0;) {
      // the reference's backends honour ctx through the HTTP request (openai.go:155); here it is polled
      // between decode chunks, so a cancel / deadline costs at most one chunk
      if (!ctx.Err().empty()) return Error::New(ctx.Err());
      const int k = left < cfg_.chunk ? left : cfg_.chunk;
      if (llmc_decode(e_, &slot, 1, k, toks.data(), nullptr) != LLMC_OK) return fail("decode");
      emit(toks.data(), (size_t)k);
      left -= k;
    }
    out->Model = req.Model;
    out->Content = content;
    out->Provider = "local";
    out->Latency = std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - start).count();
    return Error::Nil();
  }

 private:
  Error fail(const std::string& what) { return Error::New("local provider: " + what + ": " + llmc_last_error(e_)); }
  llmc_engine* e_;
  LocalConfig cfg_;
  int vocab_;
};

}  // namespace

Error NewLocal(const LocalConfig& cfg, std::shared_ptr<Provider>* out) {
  llmc_config c{};
  c.device = cfg.device;
  c.shape = cfg.shape;
  c.hidden = cfg.hidden; c.layers = cfg.layers; c.q_heads = cfg.q_heads; c.kv_heads = cfg.kv_heads;
  c.head_dim = cfg.head_dim; c.inter = cfg.inter; c.vocab = cfg.vocab; c.tied = cfg.tied;
  c.seed = cfg.seed;
  c.weights_path = cfg.weights_path.empty() ? nullptr : cfg.weights_path.c_str();
  c.max_slots = 2;
  c.max_ctx = cfg.max_ctx;
  c.flags = cfg.flags;
  llmc_engine* e = nullptr;
  if (llmc_engine_create(&c, &e) != LLMC_OK) return Error::New(std::string("local provider: ") + llmc_last_error(nullptr));
  const int vocab = cfg.shape == LLMC_SHAPE_CUSTOM ? cfg.vocab : 128256;
  *out = std::make_shared<Local>(e, cfg, vocab);
  return Error::Nil();
}

}  // namespace host
