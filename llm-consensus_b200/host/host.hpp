// host.hpp -- C++ mirror of the reference's Go packages around the provider boundary, so that a
// local model is a drop-in next to the remote ones: same type and method names, argument meaning
// and error strings as
//   internal/provider/provider.go   (Provider, Request, Response, StreamCallback, ProviderFunc)
//   internal/provider/registry.go   (Registry)
//   internal/runner/runner.go       (Runner, Callbacks, Result)
//   internal/consensus/judge.go     (Judge, the prompt template)
//   internal/output/output.go       (Result + the encoding/json behaviour main.go relies on)
// The reference is Go; no Go toolchain exists in this image, so the host
// side is C++ (the reference is compiled code) above the C ABI of include/llmc.h.  The cgo form of
// the same provider is shown in INTEGRATION.md.
#pragma once
#include <atomic>
#include <chrono>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

namespace host {

// ------------------------------------------------------------------ error / context (Go's error, context.Context)
struct Error {
  bool set = false;
  std::string msg;
  static Error Nil() { return Error{}; }
  static Error New(std::string m) { return Error{true, std::move(m)}; }
  explicit operator bool() const { return set; }
};

class Context {
 public:
  static Context Background();
  // child context that is done when the parent is, when cancel() is called, or at the deadline
  static Context WithTimeout(const Context& parent, std::chrono::nanoseconds d);
  static Context WithCancel(const Context& parent);
  void cancel() const;
  // "" while live; "context canceled" / "context deadline exceeded" once done (ctx.Err())
  std::string Err() const;
  bool Done() const { return !Err().empty(); }

 private:
  struct State {
    std::atomic<bool> cancelled{false};
    bool has_deadline = false;
    std::chrono::steady_clock::time_point deadline;
    std::shared_ptr<State> parent;
  };
  std::shared_ptr<State> st_;
};

// ------------------------------------------------------------------ provider.go
using StreamCallback = std::function<void(const std::string& chunk)>;  // provider.go:10

struct Request {  // provider.go:24-27
  std::string Model, Prompt;
};
struct Response {  // provider.go:30-35; Latency is a time.Duration (int64 ns) under the JSON key latency_ms
  std::string Model, Content, Provider;
  int64_t Latency = 0;
};

class Provider {  // provider.go:13-21
 public:
  virtual ~Provider() = default;
  virtual Error Query(const Context& ctx, const Request& req, Response* out) = 0;
  virtual Error QueryStream(const Context& ctx, const Request& req, const StreamCallback& cb, Response* out) = 0;
};

// provider.go:39-55: function adapter; QueryStream = Query + one callback with the full content
class ProviderFunc : public Provider {
 public:
  using Fn = std::function<Error(const Context&, const Request&, Response*)>;
  explicit ProviderFunc(Fn f) : f_(std::move(f)) {}
  Error Query(const Context& ctx, const Request& req, Response* out) override { return f_(ctx, req, out); }
  Error QueryStream(const Context& ctx, const Request& req, const StreamCallback& cb, Response* out) override {
    Error e = f_(ctx, req, out);
    if (e) return e;
    if (cb) cb(out->Content);
    return Error::Nil();
  }

 private:
  Fn f_;
};

// ------------------------------------------------------------------ registry.go
class Registry {
 public:
  void Register(const std::string& model, std::shared_ptr<Provider> p);
  Error Get(const std::string& model, std::shared_ptr<Provider>* out) const;  // "unknown model: %s"
  std::vector<std::string> Models() const;

 private:
  mutable std::shared_mutex mu_;
  std::map<std::string, std::shared_ptr<Provider>> providers_;
};

// ------------------------------------------------------------------ runner.go
struct Callbacks {  // runner.go:15-20
  std::function<void(const std::string& model)> OnModelStart;
  std::function<void(const std::string& model, const std::string& chunk)> OnModelStream;
  std::function<void(const std::string& model)> OnModelComplete;
  std::function<void(const std::string& model, const Error& err)> OnModelError;
};
struct RunResult {  // runner.go:23-27
  std::vector<Response> Responses;
  std::vector<std::string> Warnings, FailedModels;
};
class Runner {
 public:
  Runner(const Registry* registry, std::chrono::nanoseconds timeout) : registry_(registry), timeout_(timeout) {}
  Runner& WithCallbacks(const Callbacks* cb) {
    cb_ = cb;
    return *this;
  }
  // one thread per entry of `models` (duplicates are not deduplicated, runner.go:62); best effort;
  // responses in completion order; error "all models failed: [w1 w2]" when none succeeded
  Error Run(const Context& ctx, const std::vector<std::string>& models, const std::string& prompt, RunResult* out) const;

 private:
  const Registry* registry_;
  std::chrono::nanoseconds timeout_;
  const Callbacks* cb_ = nullptr;
};

// ------------------------------------------------------------------ judge.go
std::string RenderJudgePrompt(const std::string& prompt, const std::vector<Response>& responses);
class Judge {
 public:
  Judge(std::shared_ptr<Provider> p, std::string model) : provider_(std::move(p)), model_(std::move(model)) {}
  Error Synthesize(const Context& ctx, const std::string& prompt, const std::vector<Response>& responses,
                   std::string* out) const {
    return SynthesizeStream(ctx, prompt, responses, nullptr, out);
  }
  Error SynthesizeStream(const Context& ctx, const std::string& prompt, const std::vector<Response>& responses,
                         const StreamCallback& cb, std::string* out) const;

 private:
  std::shared_ptr<Provider> provider_;
  std::string model_;
};

// ------------------------------------------------------------------ output.go + encoding/json
struct OutputResult {  // output.go:8-15
  std::string Prompt;
  std::vector<Response> Responses;
  bool ResponsesNil = false;
  std::string Consensus, Judge;
  std::vector<std::string> Warnings, FailedModels;
};
std::string GoJSONString(const std::string& s);         // encoding/json string encoder, HTML escaping on
std::string EncodeResult(const OutputResult& r);        // SetIndent("", "  ") + Encode (trailing '\n')
std::string GoStringList(const std::vector<std::string>& v);  // fmt %v of []string

// ------------------------------------------------------------------ token codec (codec.cpp)
// ids 0x0A and 0x20..0x7E except '[' print as themselves; every other id prints as [[id]], which
// re-tokenises to the same single id.  Text bytes that are not part of such an escape map to their
// byte value.  tokenize(detokenize(ids)) == ids for all ids < vocab.
std::vector<int32_t> Tokenize(const std::string& text, int vocab);
std::string Detokenize(const int32_t* ids, size_t n);

// ------------------------------------------------------------------ local backend (local_provider.cpp)
struct LocalConfig {
  int shape = 1;             // LLMC_SHAPE_*; 0 = custom dims below
  int hidden = 0, layers = 0, q_heads = 0, kv_heads = 0, head_dim = 0, inter = 0, vocab = 128256, tied = 0;
  uint64_t seed = 0;
  int device = 0;
  std::string weights_path;  // "" = seeded random init
  int max_new_tokens = 256;
  int max_ctx = 8192;
  int chunk = 16;            // decode steps per C call; ctx is polled between chunks
  int flags = 0;
  int max_slots = 4;         // sequences that can be in flight on this instance; further callers wait for a slot
  std::vector<int32_t> stop_ids;  // generation ends before the first of these ids (e.g. 128001 / 128009 for Llama-3)
  bool prompt_cache = true;  // keep the last prompt's KV: an identical prompt is cloned, a shared prefix (>= 32 tokens) is not prefilled again
};
// provider.NewLocal: builds an engine through the C ABI (include/llmc.h); Provider tag "local"
Error NewLocal(const LocalConfig& cfg, std::shared_ptr<Provider>* out);
// forget every local instance's cached prompt (the CLI's measurement loop calls it between passes, so that a pass
// never starts with the previous pass's prompts already in the KV cache)
void DropPromptCaches();

}  // namespace host
