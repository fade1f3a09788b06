// host.cpp -- registry, runner, judge, result encoder and token codec (see host.hpp for the
// reference files each part mirrors).
#include "host.hpp"

#include <thread>

namespace host {

// ------------------------------------------------------------------ context
Context Context::Background() {
  Context c;
  c.st_ = std::make_shared<State>();
  return c;
}
Context Context::WithCancel(const Context& parent) {
  Context c;
  c.st_ = std::make_shared<State>();
  c.st_->parent = parent.st_;
  return c;
}
Context Context::WithTimeout(const Context& parent, std::chrono::nanoseconds d) {
  Context c = WithCancel(parent);
  c.st_->has_deadline = true;
  c.st_->deadline = std::chrono::steady_clock::now() + d;
  return c;
}
void Context::cancel() const {
  if (st_) st_->cancelled.store(true);
}
std::string Context::Err() const {
  const auto now = std::chrono::steady_clock::now();
  for (const State* s = st_.get(); s; s = s->parent.get()) {
    if (s->cancelled.load()) return "context canceled";
    if (s->has_deadline && now >= s->deadline) return "context deadline exceeded";
  }
  return "";
}

// ------------------------------------------------------------------ registry.go:24-53
void Registry::Register(const std::string& model, std::shared_ptr<Provider> p) {
  std::unique_lock<std::shared_mutex> lk(mu_);
  providers_[model] = std::move(p);
}
Error Registry::Get(const std::string& model, std::shared_ptr<Provider>* out) const {
  std::shared_lock<std::shared_mutex> lk(mu_);
  auto it = providers_.find(model);
  if (it == providers_.end()) return Error::New("unknown model: " + model);
  *out = it->second;
  return Error::Nil();
}
std::vector<std::string> Registry::Models() const {
  std::shared_lock<std::shared_mutex> lk(mu_);
  std::vector<std::string> v;
  for (auto& kv : providers_) v.push_back(kv.first);
  return v;
}

// ------------------------------------------------------------------ runner.go:52-131
std::string GoStringList(const std::vector<std::string>& v) {
  std::string s = "[";
  for (size_t i = 0; i < v.size(); ++i) {
    if (i) s += " ";
    s += v[i];
  }
  return s + "]";
}

Error Runner::Run(const Context& ctx, const std::vector<std::string>& models, const std::string& prompt,
                  RunResult* out) const {
  std::mutex mu;
  std::vector<Response> responses;
  std::vector<std::string> warnings, failed;
  std::vector<std::thread> threads;
  for (const std::string& model : models) {
    threads.emplace_back([&, model]() {
      Context modelCtx = Context::WithTimeout(ctx, timeout_);  // per-model timeout (runner.go:65)
      if (cb_ && cb_->OnModelStart) cb_->OnModelStart(model);
      std::shared_ptr<Provider> p;
      Error err = registry_->Get(model, &p);
      if (err) {
        {
          std::lock_guard<std::mutex> lk(mu);
          warnings.push_back(model + ": " + err.msg);
          failed.push_back(model);
        }
        if (cb_ && cb_->OnModelError) cb_->OnModelError(model, err);
        return;  // best effort: don't fail the entire run
      }
      StreamCallback streamCallback = [&](const std::string& chunk) {
        if (cb_ && cb_->OnModelStream) cb_->OnModelStream(model, chunk);
      };
      Response resp;
      err = p->QueryStream(modelCtx, Request{model, prompt}, streamCallback, &resp);
      std::lock_guard<std::mutex> lk(mu);
      if (err) {
        warnings.push_back(model + ": " + err.msg);
        failed.push_back(model);
        if (cb_ && cb_->OnModelError) cb_->OnModelError(model, err);
        return;
      }
      responses.push_back(resp);  // completion order (runner.go:109)
      if (cb_ && cb_->OnModelComplete) cb_->OnModelComplete(model);
    });
  }
  for (auto& t : threads) t.join();
  if (responses.empty()) return Error::New("all models failed: " + GoStringList(warnings));
  out->Responses = std::move(responses);
  out->Warnings = std::move(warnings);
  out->FailedModels = std::move(failed);
  return Error::Nil();
}

// ------------------------------------------------------------------ judge.go:12-105
static const char* kJudgeHead =
    "\nRole\nYou are an expert synthesis judge and careful editor. Your job is to combine multiple AI model "
    "responses into one best-possible answer to the user.\n\nInputs\nUser's original prompt:\n";
static const char* kJudgeMid = "\n\nModel responses:\n";
static const char* kJudgeTail =
    "\n\nTask\nProduce ONE final answer that directly addresses the user's original prompt by synthesizing the model "
    "responses.\n\nMethod\n1) Infer the user's intent and constraints from the original prompt (scope, tone, "
    "formatting, assumptions). Follow them.\n2) Extract the strongest points that are supported and/or repeated "
    "across responses.\n3) Resolve conflicts:\n   - Prefer statements that are more logically sound, more specific, "
    "and better justified.\n   - Prefer safer, broadly valid guidance over speculative or brittle claims.\n   - If "
    "uncertainty remains, choose the most defensible formulation and qualify it briefly.\n4) Fill gaps only when "
    "needed to make the answer complete and usable. Do not invent facts; do not add extraneous content.\n\nOutput "
    "Requirements\n- Output ONLY the final synthesized answer (no preamble, no meta-commentary, no mention of models "
    "or \xE2\x80\x9C" "consensus\xE2\x80\x9D).\n- Do not quote or reference individual model responses.\n- Keep the "
    "answer coherent, non-redundant, and well-structured (use bullets/steps/headings if helpful).\n- Match formatting "
    "appropriate to the task (e.g., code blocks for code).\n";

std::string RenderJudgePrompt(const std::string& prompt, const std::vector<Response>& responses) {
  std::string s = kJudgeHead;
  s += prompt;
  s += kJudgeMid;
  for (const Response& r : responses)  // {{range .Responses}} body, every newline literal (judge.go:21-25)
    s += "\n--- Model: " + r.Model + " | Provider: " + r.Provider + " ---\n" + r.Content + "\n\n";
  s += kJudgeTail;
  return s;
}

Error Judge::SynthesizeStream(const Context& ctx, const std::string& prompt, const std::vector<Response>& responses,
                              const StreamCallback& cb, std::string* out) const {
  if (responses.empty()) return Error::New("no responses to synthesize");
  if (responses.size() == 1) {  // no consensus needed; the judge provider is NOT called (judge.go:74-79)
    if (cb) cb(responses[0].Content);
    *out = responses[0].Content;
    return Error::Nil();
  }
  Response resp;
  Error err = provider_->QueryStream(ctx, Request{model_, RenderJudgePrompt(prompt, responses)}, cb, &resp);
  if (err) return Error::New("judge query failed: " + err.msg);
  *out = resp.Content;
  return Error::Nil();
}

// ------------------------------------------------------------------ encoding/json
// utf8.DecodeRune: returns the sequence length and code point, or (1, -1) for an invalid byte
static int decode_rune(const unsigned char* b, size_t n, int* cp) {
  const unsigned c = b[0];
  int need;
  unsigned lo = 0x80, hi = 0xBF;
  int v;
  if (c >= 0xC2 && c <= 0xDF) { need = 1; v = c & 0x1F; }
  else if (c == 0xE0) { need = 2; lo = 0xA0; v = c & 0x0F; }
  else if ((c >= 0xE1 && c <= 0xEC) || c == 0xEE || c == 0xEF) { need = 2; v = c & 0x0F; }
  else if (c == 0xED) { need = 2; hi = 0x9F; v = c & 0x0F; }
  else if (c == 0xF0) { need = 3; lo = 0x90; v = c & 0x07; }
  else if (c >= 0xF1 && c <= 0xF3) { need = 3; v = c & 0x07; }
  else if (c == 0xF4) { need = 3; hi = 0x8F; v = c & 0x07; }
  else { *cp = -1; return 1; }
  for (int k = 1; k <= need; ++k) {
    if ((size_t)k >= n) { *cp = -1; return 1; }
    const unsigned x = b[k];
    const unsigned l = k == 1 ? lo : 0x80, h = k == 1 ? hi : 0xBF;
    if (x < l || x > h) { *cp = -1; return 1; }
    v = (v << 6) | (x & 0x3F);
  }
  *cp = v;
  return need + 1;
}

std::string GoJSONString(const std::string& s) {
  static const char* hex = "0123456789abcdef";
  std::string o = "\"";
  const unsigned char* b = reinterpret_cast<const unsigned char*>(s.data());
  const size_t n = s.size();
  for (size_t i = 0; i < n;) {
    const unsigned c = b[i];
    if (c < 0x80) {
      switch (c) {
        case '"': o += "\\\""; break;
        case '\\': o += "\\\\"; break;
        case '\n': o += "\\n"; break;
        case '\r': o += "\\r"; break;
        case '\t': o += "\\t"; break;
        case '\b': o += "\\b"; break;  // go >= 1.22
        case '\f': o += "\\f"; break;  // go >= 1.22
        default:
          if (c < 0x20 || c == '<' || c == '>' || c == '&') {
            o += "\\u00";
            o += hex[c >> 4];
            o += hex[c & 0xF];
          } else {
            o += (char)c;
          }
      }
      ++i;
      continue;
    }
    int cp;
    const int len = decode_rune(b + i, n - i, &cp);
    if (cp < 0) {
      o += "\\ufffd";
      ++i;
      continue;
    }
    if (cp == 0x2028 || cp == 0x2029) {
      o += "\\u202";
      o += hex[cp & 0xF];
    } else {
      o.append(s, i, len);
    }
    i += len;
  }
  return o + "\"";
}

std::string EncodeResult(const OutputResult& r) {
  std::string o = "{\n  \"prompt\": " + GoJSONString(r.Prompt) + ",\n";
  if (r.ResponsesNil) {
    o += "  \"responses\": null,\n";
  } else if (r.Responses.empty()) {
    o += "  \"responses\": [],\n";
  } else {
    o += "  \"responses\": [\n";
    for (size_t k = 0; k < r.Responses.size(); ++k) {
      const Response& x = r.Responses[k];
      o += "    {\n      \"model\": " + GoJSONString(x.Model) + ",\n      \"content\": " + GoJSONString(x.Content) +
           ",\n      \"provider\": " + GoJSONString(x.Provider) + ",\n      \"latency_ms\": " +
           std::to_string((long long)x.Latency) + "\n    }" + (k + 1 < r.Responses.size() ? "," : "") + "\n";
    }
    o += "  ],\n";
  }
  o += "  \"consensus\": " + GoJSONString(r.Consensus) + ",\n  \"judge\": " + GoJSONString(r.Judge);
  auto list = [&](const char* key, const std::vector<std::string>& v) {
    if (v.empty()) return;  // omitempty
    o += std::string(",\n  \"") + key + "\": [\n";
    for (size_t k = 0; k < v.size(); ++k) o += "    " + GoJSONString(v[k]) + (k + 1 < v.size() ? "," : "") + "\n";
    o += "  ]";
  };
  list("warnings", r.Warnings);
  list("failed_models", r.FailedModels);
  return o + "\n}\n";
}

// ------------------------------------------------------------------ token codec
static inline bool plain_id(int t) { return t == 0x0A || (t >= 0x20 && t <= 0x7E && t != '['); }

std::string Detokenize(const int32_t* ids, size_t n) {
  std::string s;
  for (size_t i = 0; i < n; ++i) {
    if (plain_id(ids[i])) s += (char)ids[i];
    else s += "[[" + std::to_string(ids[i]) + "]]";
  }
  return s;
}

std::vector<int32_t> Tokenize(const std::string& text, int vocab) {
  std::vector<int32_t> ids;
  const unsigned char* b = reinterpret_cast<const unsigned char*>(text.data());
  const size_t n = text.size();
  for (size_t i = 0; i < n;) {
    if (b[i] == '[' && i + 1 < n && b[i + 1] == '[') {
      size_t j = i + 2;
      while (j < n && j - (i + 2) < 7 && b[j] >= '0' && b[j] <= '9') ++j;
      const size_t nd = j - (i + 2);
      if (nd >= 1 && j + 1 < n && b[j] == ']' && b[j + 1] == ']') {
        long v = 0;
        for (size_t k = i + 2; k < j; ++k) v = v * 10 + (b[k] - '0');
        if ((nd == 1 || b[i + 2] != '0') && v < vocab && !plain_id((int)v)) {  // canonical escapes only
          ids.push_back((int32_t)v);
          i = j + 2;
          continue;
        }
      }
    }
    ids.push_back(b[i]);
    ++i;
  }
  return ids;
}

}  // namespace host
