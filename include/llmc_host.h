/* llmc_host.h -- plain-C test hooks over the C++ host mirror of the reference's orchestration
 * (llm-consensus_b200/host), exported by libllmc_host.so.  The product entry point of the host
 * side is the `llm-consensus` binary (host/main.cpp); these hooks only let the Python tests
 * compare its pieces with oracle/orchestration.py byte for byte.
 *
 *   llmc_host_render_judge_prompt   text/template of /root/reference/internal/consensus/judge.go:12-44
 *   llmc_host_encode_result         output.Result through encoding/json as main.go:227-229 sets it up
 *                                   (/root/reference/internal/output/output.go:8-15)
 *   llmc_host_run_stub_scenario     runner.Run + Judge.Synthesize over in-process stub providers
 *                                   (/root/reference/internal/runner/runner.go:52-131, judge.go:68-105),
 *                                   the shape of the reference's own table tests (runner_test.go:12-129)
 *   llmc_host_tokenize/detokenize   the local provider's reversible text<->id codec (no reference counterpart:
 *                                   the reference deals in strings only)
 *
 * Every function returns the length of the result; it is copied (NUL-terminated) when cap suffices. */
#ifndef LLMC_HOST_H
#define LLMC_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int64_t llmc_host_render_judge_prompt(const char* prompt, int n, const char* const* models,
                                      const char* const* providers, const char* const* contents, char* out,
                                      int64_t cap);
int64_t llmc_host_encode_result(const char* prompt, int64_t prompt_len, int n, const char* const* models,
                                const char* const* providers, const char* const* contents, const int64_t* latencies,
                                const char* consensus, const char* judge, int n_warn, const char* const* warnings,
                                int n_fail, const char* const* failed, char* out, int64_t cap);
int64_t llmc_host_tokenize(const char* text, int64_t len, int vocab, int32_t* out, int64_t cap);
int64_t llmc_host_detokenize(const int32_t* ids, int64_t n, char* out, int64_t cap);
int64_t llmc_host_run_stub_scenario(const char* prompt, int n, const char* const* models, const char* const* providers,
                                    const char* const* contents, const int* behaviour, const int* delay_ms,
                                    int timeout_ms, const char* judge_model, const char* judge_content, int judge_fail,
                                    char* out, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif
