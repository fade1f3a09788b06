// main.cpp -- C++ mirror of cmd/llm-consensus/main.go with the local backend wired in: the same
// flags (--models --judge --file --output --data-dir --timeout --quiet/-q --json --no-save
// --version, main.go:312-322), prompt sources (main.go:363-393), registry construction
// (main.go:395-438), runner -> judge -> output flow (main.go:132-184) and output branches
// (main.go:187-273), plus the flags BASELINE.json's north star adds for the local provider:
// --weights, --seed, --device (and --max-new-tokens).  The ANSI progress UI (internal/ui) is out of
// scope: stderr stays quiet, which is what the reference does when stderr is not a terminal
// (main.go:95).  Remote model names are recognised but their HTTP backends are not part of this
// library; they fail with the same "API key not set" class of error a user would see without keys.
#include <signal.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <random>
#include <sstream>
#include <thread>

#include "../../include/llmc.h"
#include "host.hpp"

using namespace host;

namespace {

const char* kDefaultJudge = "gpt-5.2-pro-2025-12-11";  // main.go:34
host::Context g_root = host::Context::Background();

struct Config {
  std::vector<std::string> models;
  std::string judge = kDefaultJudge, file, output, dataDir = "data", prompt, weights;
  int timeout = 120, device = 0, maxNew = 256, maxCtx = 8192;
  uint64_t seed = 0;
  bool quiet = false, json = false, noSave = false, version = false;
  std::vector<std::string> stubs;  // NAME=PROVIDER:DELAY_MS:CONTENT  (test stand-ins for remote providers)
};

enum ProviderType { ProviderOpenAI, ProviderAnthropic, ProviderGoogle, ProviderLocal };
const std::map<std::string, ProviderType>& knownModels() {  // main.go:49-61 + local names
  static const std::map<std::string, ProviderType> m = {
      {"gpt-5.2-2025-12-11", ProviderOpenAI},   {"gpt-5.2-pro-2025-12-11", ProviderOpenAI},
      {"claude-sonnet-4-5", ProviderAnthropic}, {"claude-haiku-4-5", ProviderAnthropic},
      {"claude-opus-4-5", ProviderAnthropic},   {"gemini-3-pro-preview", ProviderGoogle},
      {"local-llama-1b", ProviderLocal},        {"local-llama-8b", ProviderLocal},
      {"local-tiny", ProviderLocal}};
  return m;
}

std::string trim(const std::string& s) {
  size_t a = 0, b = s.size();
  while (a < b && isspace((unsigned char)s[a])) ++a;
  while (b > a && isspace((unsigned char)s[b - 1])) --b;
  return s.substr(a, b - a);
}

// Go's flag package: -name / --name, "-name=value" or "-name value"; parsing stops at the first non-flag
Error parseFlags(int argc, char** argv, Config* cfg) {
  std::string modelsStr;
  std::vector<std::string> args;
  int i = 1;
  for (; i < argc; ++i) {
    std::string a = argv[i];
    if (a == "--") { ++i; break; }
    if (a.size() < 2 || a[0] != '-') break;
    std::string name = a.substr(a[1] == '-' ? 2 : 1), value;
    bool has_value = false;
    const size_t eq = name.find('=');
    if (eq != std::string::npos) {
      value = name.substr(eq + 1);
      name = name.substr(0, eq);
      has_value = true;
    }
    auto boolean = [&](bool* dst) {
      if (!has_value) { *dst = true; return Error::Nil(); }
      if (value == "true" || value == "1" || value == "t" || value == "T" || value == "TRUE" || value == "True") *dst = true;
      else if (value == "false" || value == "0" || value == "f" || value == "F" || value == "FALSE" || value == "False") *dst = false;
      else return Error::New("invalid boolean value \"" + value + "\" for -" + name + ": parse error");
      return Error::Nil();
    };
    auto need = [&]() -> Error {
      if (has_value) return Error::Nil();
      if (i + 1 >= argc) return Error::New("flag needs an argument: -" + name);
      value = argv[++i];
      return Error::Nil();
    };
    auto integer = [&](long long* dst) -> Error {
      if (Error e = need()) return e;
      char* end = nullptr;
      *dst = strtoll(value.c_str(), &end, 10);
      if (value.empty() || *end) return Error::New("invalid value \"" + value + "\" for flag -" + name + ": parse error");
      return Error::Nil();
    };
    Error e = Error::Nil();
    long long n = 0;
    if (name == "quiet" || name == "q") e = boolean(&cfg->quiet);
    else if (name == "json") e = boolean(&cfg->json);
    else if (name == "no-save") e = boolean(&cfg->noSave);
    else if (name == "version") e = boolean(&cfg->version);
    else if (name == "models") { e = need(); modelsStr = value; }
    else if (name == "judge") { e = need(); cfg->judge = value; }
    else if (name == "file") { e = need(); cfg->file = value; }
    else if (name == "output") { e = need(); cfg->output = value; }
    else if (name == "data-dir") { e = need(); cfg->dataDir = value; }
    else if (name == "weights") { e = need(); cfg->weights = value; }
    else if (name == "stub") { e = need(); cfg->stubs.push_back(value); }
    else if (name == "timeout") { e = integer(&n); cfg->timeout = (int)n; }
    else if (name == "device") { e = integer(&n); cfg->device = (int)n; }
    else if (name == "max-new-tokens") { e = integer(&n); cfg->maxNew = (int)n; }
    else if (name == "max-ctx") { e = integer(&n); cfg->maxCtx = (int)n; }
    else if (name == "seed") { e = integer(&n); cfg->seed = (uint64_t)n; }
    else e = Error::New("flag provided but not defined: -" + name);
    if (e) return e;
  }
  for (; i < argc; ++i) args.push_back(argv[i]);
  if (cfg->version) return Error::Nil();
  if (modelsStr.empty()) return Error::New("--models flag is required");
  std::stringstream ss(modelsStr);
  std::string item;
  while (std::getline(ss, item, ',')) cfg->models.push_back(trim(item));
  if (!modelsStr.empty() && modelsStr.back() == ',') cfg->models.push_back("");  // strings.Split keeps the tail
  // prompt: positional arg > --file > stdin (main.go:363-393)
  if (!args.empty()) {
    for (size_t k = 0; k < args.size(); ++k) cfg->prompt += (k ? " " : "") + args[k];
    return Error::Nil();
  }
  if (!cfg->file.empty()) {
    std::ifstream f(cfg->file, std::ios::binary);
    if (!f) return Error::New("reading prompt file: open " + cfg->file + ": no such file or directory");
    std::stringstream buf;
    buf << f.rdbuf();
    cfg->prompt = trim(buf.str());
    return Error::Nil();
  }
  struct stat st;
  if (fstat(0, &st) == 0 && !S_ISCHR(st.st_mode)) {
    std::string line;
    bool first = true;
    while (std::getline(std::cin, line)) {
      if (!line.empty() && line.back() == '\r') line.pop_back();  // bufio.ScanLines drops a trailing \r
      cfg->prompt += (first ? "" : "\n") + line;
      first = false;
    }
    return Error::Nil();
  }
  return Error::New("no prompt provided: use positional argument, --file, or pipe to stdin");
}

// "local-llama-1b-s3" -> base name + seed offset 3
bool splitSeedSuffix(const std::string& model, std::string* base, uint64_t* off) {
  const size_t p = model.rfind("-s");
  *base = model;
  *off = 0;
  if (p == std::string::npos || p + 2 >= model.size()) return false;
  for (size_t k = p + 2; k < model.size(); ++k)
    if (!isdigit((unsigned char)model[k])) return false;
  *base = model.substr(0, p);
  *off = strtoull(model.c_str() + p + 2, nullptr, 10);
  return true;
}

// uses: how many QueryStream calls can be in flight on this model at once (its entries in --models, duplicates
// included as runner.go:62 starts one goroutine per entry, plus one if it is also the judge)
Error createProvider(const std::string& model, const Config& cfg, int* next_device, int uses,
                     std::shared_ptr<Provider>* out) {
  for (const std::string& s : cfg.stubs) {  // NAME=PROVIDER:DELAY_MS:CONTENT
    const size_t eq = s.find('=');
    if (eq == std::string::npos || s.substr(0, eq) != model) continue;
    const size_t c1 = s.find(':', eq + 1), c2 = c1 == std::string::npos ? c1 : s.find(':', c1 + 1);
    if (c2 == std::string::npos) return Error::New("bad --stub spec: " + s);
    const std::string tag = s.substr(eq + 1, c1 - eq - 1), content = s.substr(c2 + 1);
    const int delay = atoi(s.substr(c1 + 1, c2 - c1 - 1).c_str());
    *out = std::make_shared<ProviderFunc>([tag, content, delay](const Context& ctx, const Request& req, Response* r) {
      const auto start = std::chrono::steady_clock::now();
      const bool hang = content == "!hang";
      for (int waited = 0; hang || waited < delay; ++waited) {
        if (!ctx.Err().empty()) return Error::New(ctx.Err());
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
      }
      if (content.rfind("!error:", 0) == 0) return Error::New(content.substr(7));
      r->Model = req.Model;
      r->Content = content == "!echo" ? req.Prompt : content;
      r->Provider = tag;
      r->Latency = std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - start).count();
      return Error::Nil();
    });
    return Error::Nil();
  }
  std::string base;
  uint64_t off = 0;
  splitSeedSuffix(model, &base, &off);
  auto it = knownModels().find(model);
  if (it == knownModels().end()) it = knownModels().find(base);
  if (it == knownModels().end() || (it->second != ProviderLocal && base != model && knownModels().count(model) == 0)) {
    std::vector<std::string> available;
    for (auto& kv : knownModels()) available.push_back(kv.first);  // sorted; the Go map order is random
    return Error::New("unknown model \"" + model + "\"; available models: " + GoStringList(available));
  }
  switch (it->second) {
    case ProviderOpenAI: return Error::New("OPENAI_API_KEY environment variable not set (remote backends are outside this library)");
    case ProviderAnthropic: return Error::New("ANTHROPIC_API_KEY environment variable not set (remote backends are outside this library)");
    case ProviderGoogle: return Error::New("GOOGLE_API_KEY environment variable not set (remote backends are outside this library)");
    case ProviderLocal: {
      LocalConfig lc;
      if (it->first == "local-llama-1b") lc.shape = LLMC_SHAPE_LLAMA_1B;
      else if (it->first == "local-llama-8b") lc.shape = LLMC_SHAPE_LLAMA_8B;
      else {
        lc.shape = LLMC_SHAPE_CUSTOM;
        lc.hidden = 256; lc.layers = 2; lc.q_heads = 8; lc.kv_heads = 2; lc.head_dim = 32; lc.inter = 512;
        lc.vocab = 1024; lc.tied = 1;
      }
      lc.seed = cfg.seed + off;
      const int ndev = llmc_device_count();
      lc.device = ndev > 0 ? (cfg.device + (*next_device)++) % ndev : cfg.device;  // one instance per GPU, round robin
      lc.weights_path = cfg.weights;
      lc.max_new_tokens = cfg.maxNew;
      lc.max_ctx = cfg.maxCtx;
      lc.max_slots = std::max(2, uses);
      return NewLocal(lc, out);
    }
  }
  return Error::New("unhandled provider type for model " + model);
}

Error initRegistry(const Config& cfg, Registry* reg) {  // main.go:395-415
  std::vector<std::string> needed;  // first-appearance order: model i -> GPU i, the judge last
  auto add = [&](const std::string& m) {
    if (std::find(needed.begin(), needed.end(), m) == needed.end()) needed.push_back(m);
  };
  for (auto& m : cfg.models) add(m);
  add(cfg.judge);
  // Device slots are dealt in first-appearance order exactly as a sequential loop would (createProvider advances the
  // counter only for local models); the engines themselves -- CUDA context, weights, KV pool: about a second each --
  // are then built concurrently, one thread per model.  The first error in list order is reported, like main.go:406-412.
  if (getenv("LLMC_INIT_TRACE")) {   // how long the CUDA driver itself takes to come up (all GPUs of the node)
    const auto t0 = std::chrono::steady_clock::now();
    const int n = llmc_device_count();
    fprintf(stderr, "[init] CUDA driver up, %d device(s): %.1f ms\n", n,
            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  }
  int next_device = 0;
  std::vector<int> slot(needed.size(), 0);
  for (size_t i = 0; i < needed.size(); ++i) {
    const std::string& model = needed[i];
    bool stubbed = false;
    for (const std::string& st : cfg.stubs) {
      const size_t eq = st.find('=');
      if (eq != std::string::npos && st.substr(0, eq) == model) stubbed = true;
    }
    std::string base;
    uint64_t off = 0;
    splitSeedSuffix(model, &base, &off);
    auto it = knownModels().find(model);
    if (it == knownModels().end()) it = knownModels().find(base);
    const bool local = !stubbed && it != knownModels().end() && it->second == ProviderLocal;
    // the judge runs on GPU 0 (BASELINE.json north star) unless it is also one of the fan-out models
    const bool judge_only = model == cfg.judge && std::find(cfg.models.begin(), cfg.models.end(), model) == cfg.models.end();
    slot[i] = (local && !judge_only) ? next_device++ : 0;
  }
  std::vector<std::shared_ptr<Provider>> made(needed.size());
  std::vector<Error> errs(needed.size(), Error::Nil());
  std::vector<std::thread> workers;
  const bool serial = getenv("LLMC_SERIAL_INIT") != nullptr;   // measurement aid: build the engines one after another
  for (size_t i = 0; i < needed.size(); ++i) {
    workers.emplace_back([&, i] {
      int dev = slot[i];
      const int uses = (int)std::count(cfg.models.begin(), cfg.models.end(), needed[i]) + (needed[i] == cfg.judge ? 1 : 0);
      errs[i] = createProvider(needed[i], cfg, &dev, uses, &made[i]);
    });
    if (serial) workers.back().join();
  }
  for (auto& w : workers)
    if (w.joinable()) w.join();
  for (size_t i = 0; i < needed.size(); ++i) {
    if (errs[i]) return Error::New("initializing provider for " + needed[i] + ": " + errs[i].msg);
    reg->Register(needed[i], made[i]);
  }
  return Error::Nil();
}

std::string generateRunID() {  // main.go:280-285
  char ts[32];
  time_t t = time(nullptr);
  struct tm tmv;
  localtime_r(&t, &tmv);
  strftime(ts, sizeof ts, "%Y%m%d-%H%M%S", &tmv);
  std::random_device rd;
  char suf[8];
  snprintf(suf, sizeof suf, "%02x%02x%02x", rd() & 0xFF, rd() & 0xFF, rd() & 0xFF);
  return std::string(ts) + "-" + suf;
}

Error writeFile(const std::string& path, const std::string& data) {
  std::ofstream f(path, std::ios::binary);
  if (!f) return Error::New("open " + path + ": " + strerror(errno));
  f << data;
  return Error::Nil();
}

Error run(int argc, char** argv) {
  Config cfg;
  if (Error e = parseFlags(argc, argv, &cfg)) return e;
  if (cfg.version) {
    printf("llm-consensus dev (B200 local provider, %s)\n  commit: none\n  built:  unknown\n", llmc_version());
    return Error::Nil();
  }
  using clk = std::chrono::steady_clock;
  auto ms_since = [](clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); };
  // LLMC_BENCH_REPEAT=N (measurement aid, not a reference flag): run the fan-out + judge N times on the same
  // providers and report each pass on stderr; the output written is the last pass's
  const char* rep_env = getenv("LLMC_BENCH_REPEAT");
  const int repeat = rep_env ? std::max(1, atoi(rep_env)) : 1;
  const auto t_init = clk::now();
  Registry registry;
  if (Error e = initRegistry(cfg, &registry)) return e;
  if (rep_env) fprintf(stderr, "[bench] providers ready in %.1f ms\n", ms_since(t_init));
  Runner r(&registry, std::chrono::seconds(cfg.timeout));
  RunResult result;
  std::string consensus;
  for (int pass = 0; pass < repeat; ++pass) {
    if (pass > 0) DropPromptCaches();   // every pass starts like a fresh process: nothing of the last one is reused
    result = RunResult();
    consensus.clear();
    const auto t0 = clk::now();
    if (Error e = r.Run(g_root, cfg.models, cfg.prompt, &result)) return Error::New("running queries: " + e.msg);
    const double fan_ms = ms_since(t0);
    std::shared_ptr<Provider> judgeProvider;
    if (Error e = registry.Get(cfg.judge, &judgeProvider)) return Error::New("judge model " + cfg.judge + ": " + e.msg);
    Judge judge(judgeProvider, cfg.judge);
    const auto t1 = clk::now();
    if (Error e = judge.SynthesizeStream(g_root, cfg.prompt, result.Responses, nullptr, &consensus))
      return Error::New("consensus synthesis: " + e.msg);
    if (rep_env)
      fprintf(stderr, "[bench] pass %d: fan-out %.1f ms (%zu responses, %zu failed), judge %.1f ms, run %.1f ms\n", pass,
              fan_ms, result.Responses.size(), result.FailedModels.size(), ms_since(t1), ms_since(t0));
  }
  OutputResult out;
  out.Prompt = cfg.prompt;
  out.Responses = result.Responses;
  out.Consensus = consensus;
  out.Judge = cfg.judge;
  out.Warnings = result.Warnings;
  out.FailedModels = result.FailedModels;
  std::string outputPath;
  if (!cfg.output.empty()) {
    outputPath = cfg.output;
  } else if (!cfg.json && !cfg.noSave) {  // auto-save to data/<run-id>/ (main.go:191-216)
    const std::string runDir = cfg.dataDir + "/" + generateRunID();
    if (mkdir(cfg.dataDir.c_str(), 0755) != 0 && errno != EEXIST)
      return Error::New("creating run directory: mkdir " + cfg.dataDir + ": " + strerror(errno));
    if (mkdir(runDir.c_str(), 0755) != 0 && errno != EEXIST)
      return Error::New("creating run directory: mkdir " + runDir + ": " + strerror(errno));
    outputPath = runDir + "/result.json";
    writeFile(runDir + "/prompt.txt", cfg.prompt);      // failures here are non-fatal in the reference
    writeFile(runDir + "/consensus.md", consensus);
  }
  const std::string doc = EncodeResult(out);
  if (!outputPath.empty()) {
    if (Error e = writeFile(outputPath, doc)) return Error::New("creating output file: " + e.msg);
  } else {
    fputs(doc.c_str(), stdout);  // --json, or non-interactive: JSON to stdout
  }
  return Error::Nil();
}

void onSignal(int) { g_root.cancel(); }

}  // namespace

int main(int argc, char** argv) {
  signal(SIGINT, onSignal);   // signal.NotifyContext(ctx, SIGINT, SIGTERM), main.go:90
  signal(SIGTERM, onSignal);
  if (Error e = run(argc, argv)) {
    fprintf(stderr, "error: %s\n", e.msg.c_str());  // main.go:77-80
    return 1;
  }
  return 0;
}
