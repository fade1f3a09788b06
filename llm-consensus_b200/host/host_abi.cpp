// host_abi.cpp -- plain-C test hooks over the host mirror (include/llmc_host.h) so that the Python
// tests can compare its template rendering, JSON encoding, codec and runner/judge semantics with
// the orchestration oracle without spawning the CLI.
#include <cstring>
#include <thread>

#include "../../include/llmc_host.h"
#include "host.hpp"

using namespace host;

namespace {
int copy_out(const std::string& s, char* out, int64_t cap) {
  if (out && cap > (int64_t)s.size()) {
    memcpy(out, s.data(), s.size());
    out[s.size()] = 0;
  }
  return (int)s.size();
}
std::vector<Response> responses_from(int n, const char* const* models, const char* const* providers,
                                     const char* const* contents, const int64_t* latencies) {
  std::vector<Response> v(n);
  for (int i = 0; i < n; ++i) {
    v[i].Model = models[i];
    v[i].Provider = providers[i];
    v[i].Content = contents[i];
    v[i].Latency = latencies ? latencies[i] : 0;
  }
  return v;
}
}  // namespace

extern "C" {

int64_t llmc_host_render_judge_prompt(const char* prompt, int n, const char* const* models,
                                      const char* const* providers, const char* const* contents, char* out,
                                      int64_t cap) {
  return copy_out(RenderJudgePrompt(prompt, responses_from(n, models, providers, contents, nullptr)), out, cap);
}

int64_t llmc_host_encode_result(const char* prompt, int64_t prompt_len, int n, const char* const* models,
                                const char* const* providers, const char* const* contents, const int64_t* latencies,
                                const char* consensus, const char* judge, int n_warn, const char* const* warnings,
                                int n_fail, const char* const* failed, char* out, int64_t cap) {
  OutputResult r;
  r.Prompt.assign(prompt, (size_t)prompt_len);
  r.Responses = responses_from(n, models, providers, contents, latencies);
  r.Consensus = consensus;
  r.Judge = judge;
  for (int i = 0; i < n_warn; ++i) r.Warnings.push_back(warnings[i]);
  for (int i = 0; i < n_fail; ++i) r.FailedModels.push_back(failed[i]);
  return copy_out(EncodeResult(r), out, cap);
}

int64_t llmc_host_tokenize(const char* text, int64_t len, int vocab, int32_t* out, int64_t cap) {
  std::vector<int32_t> ids = Tokenize(std::string(text, (size_t)len), vocab);
  if (out && cap >= (int64_t)ids.size()) memcpy(out, ids.data(), ids.size() * sizeof(int32_t));
  return (int64_t)ids.size();
}

int64_t llmc_host_detokenize(const int32_t* ids, int64_t n, char* out, int64_t cap) {
  return copy_out(Detokenize(ids, (size_t)n), out, cap);
}

// Runs Runner.Run + Judge.Synthesize over stub providers described by parallel arrays:
//   behaviour[i]: 0 = succeed with contents[i] after delay_ms[i]; 1 = fail with contents[i] as the error;
//                 2 = block until the context is done; 3 = not registered at all.
// The judge stub answers judge_content (or fails when judge_fail != 0).  Writes result.json (or
// "error: <msg>\n") to out; returns its length.
int64_t llmc_host_run_stub_scenario(const char* prompt, int n, const char* const* models, const char* const* providers,
                                    const char* const* contents, const int* behaviour, const int* delay_ms,
                                    int timeout_ms, const char* judge_model, const char* judge_content, int judge_fail,
                                    char* out, int64_t cap) {
  Registry reg;
  for (int i = 0; i < n; ++i) {
    if (behaviour[i] == 3) continue;
    const std::string tag = providers[i], content = contents[i];
    const int beh = behaviour[i], delay = delay_ms[i];
    reg.Register(models[i], std::make_shared<ProviderFunc>([=](const Context& ctx, const Request& req, Response* r) {
      for (int waited = 0; beh == 2 || waited < delay; ++waited) {
        if (!ctx.Err().empty()) return Error::New(ctx.Err());
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
      }
      if (beh == 1) return Error::New(content);
      r->Model = req.Model;
      r->Content = content;
      r->Provider = tag;
      r->Latency = 1000000LL * delay;
      return Error::Nil();
    }));
  }
  std::string judge_prompt_seen;
  const std::string jc = judge_content;
  auto judge_p = std::make_shared<ProviderFunc>([&, jc](const Context&, const Request& req, Response* r) {
    judge_prompt_seen = req.Prompt;
    if (judge_fail) return Error::New(jc);
    r->Content = jc;
    return Error::Nil();
  });
  Context root = Context::Background();
  std::vector<std::string> names;
  for (int i = 0; i < n; ++i) names.push_back(models[i]);
  Runner runner(&reg, std::chrono::milliseconds(timeout_ms));
  RunResult rr;
  if (Error e = runner.Run(root, names, prompt, &rr)) return copy_out("error: running queries: " + e.msg + "\n", out, cap);
  Judge judge(judge_p, judge_model);
  std::string consensus;
  if (Error e = judge.Synthesize(root, prompt, rr.Responses, &consensus))
    return copy_out("error: consensus synthesis: " + e.msg + "\n", out, cap);
  OutputResult o;
  o.Prompt = prompt;
  o.Responses = rr.Responses;
  o.Consensus = consensus;
  o.Judge = judge_model;
  o.Warnings = rr.Warnings;
  o.FailedModels = rr.FailedModels;
  return copy_out(EncodeResult(o), out, cap);
}

}  // extern "C"
