"""CPU restatement of the reference's orchestration around the provider boundary: the runner
fan-out, the judge prompt template and the result.json encoder.

TEST INFRASTRUCTURE ONLY (see oracle/ref.py).  It is what tests compare the C++ host driver
(llm-consensus_b200/host) with, byte for byte.

Follows, function by function:
  * render_judge_prompt   /root/reference/internal/consensus/judge.go:12-44 (template text), :82-93
  * synthesize            /root/reference/internal/consensus/judge.go:68-105
  * run_models            /root/reference/internal/runner/runner.go:52-131
  * encode_result         /root/reference/internal/output/output.go:8-15,
                          /root/reference/internal/provider/provider.go:30-35 and the encoder
                          set-up in /root/reference/cmd/llm-consensus/main.go:227-229, plus the
                          published behaviour of Go's encoding/json (go1.24, go.mod:5): field order
                          as declared, SetIndent("", "  "), omitempty on warnings/failed_models,
                          HTML escaping on, invalid UTF-8 -> \\ufffd, time.Duration -> int64 ns,
                          trailing newline.
  * go_string_list        fmt's %v of a []string ("[a b]"), used by runner.go:123

PARITY: UNPINNED BY A RUN OF THE REFERENCE -- Go is not installed in this image, so the
reference cannot be executed; the reference's own tests (runner_test.go:12-129,
judge_test.go:13-136) only pin counts and substrings, which tests/test_orchestration.py
checks against this file.  The 1437-byte / SHA-256 golden of SURVEY.md section 8(c) was
produced by an earlier restatement of the same template and is re-checked there.
"""
import threading
import time

JUDGE_TEMPLATE_HEAD = """
Role
You are an expert synthesis judge and careful editor. Your job is to combine multiple AI model responses into one best-possible answer to the user.

Inputs
User's original prompt:
"""
JUDGE_TEMPLATE_MID = """

Model responses:
"""
JUDGE_TEMPLATE_TAIL = """

Task
Produce ONE final answer that directly addresses the user's original prompt by synthesizing the model responses.

Method
1) Infer the user's intent and constraints from the original prompt (scope, tone, formatting, assumptions). Follow them.
2) Extract the strongest points that are supported and/or repeated across responses.
3) Resolve conflicts:
   - Prefer statements that are more logically sound, more specific, and better justified.
   - Prefer safer, broadly valid guidance over speculative or brittle claims.
   - If uncertainty remains, choose the most defensible formulation and qualify it briefly.
4) Fill gaps only when needed to make the answer complete and usable. Do not invent facts; do not add extraneous content.

Output Requirements
- Output ONLY the final synthesized answer (no preamble, no meta-commentary, no mention of models or “consensus”).
- Do not quote or reference individual model responses.
- Keep the answer coherent, non-redundant, and well-structured (use bullets/steps/headings if helpful).
- Match formatting appropriate to the task (e.g., code blocks for code).
"""


def render_judge_prompt(prompt, responses):
    """text/template execution of judge.go:12-44: no escaping, every newline literal.
    responses: list of dicts with model, provider, content."""
    out = [JUDGE_TEMPLATE_HEAD, prompt, JUDGE_TEMPLATE_MID]
    for r in responses:
        # "{{range .Responses}}\n--- Model: M | Provider: P ---\nCONTENT\n\n{{end}}"
        out.append("\n--- Model: %s | Provider: %s ---\n%s\n\n" % (r["model"], r["provider"], r["content"]))
    out.append(JUDGE_TEMPLATE_TAIL)
    return "".join(out)


class ProviderError(Exception):
    pass


def synthesize(judge_fn, judge_model, prompt, responses, callback=None):
    """judge.go:68-105.  judge_fn(model, prompt, callback) -> content or raises ProviderError."""
    if len(responses) == 0:
        raise ProviderError("no responses to synthesize")
    if len(responses) == 1:
        if callback:
            callback(responses[0]["content"])
        return responses[0]["content"]
    rendered = render_judge_prompt(prompt, responses)
    try:
        return judge_fn(judge_model, rendered, callback)
    except ProviderError as e:
        raise ProviderError("judge query failed: %s" % e)


def go_string_list(items):
    return "[" + " ".join(items) + "]"


def run_models(registry, models, prompt, timeout_s=120.0):
    """runner.go:52-131.  registry: dict model -> fn(model, prompt, deadline) returning a response dict
    or raising ProviderError.  Responses are appended in COMPLETION order (runner.go:109)."""
    lock = threading.Lock()
    responses, warnings, failed = [], [], []

    def work(model):
        deadline = time.monotonic() + timeout_s
        fn = registry.get(model)
        if fn is None:
            with lock:
                warnings.append("%s: unknown model: %s" % (model, model))
                failed.append(model)
            return
        try:
            resp = fn(model, prompt, deadline)
        except ProviderError as e:
            with lock:
                warnings.append("%s: %s" % (model, e))
                failed.append(model)
            return
        with lock:
            responses.append(resp)

    threads = [threading.Thread(target=work, args=(m,)) for m in models]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if not responses:
        raise ProviderError("all models failed: " + go_string_list(warnings))
    return {"responses": responses, "warnings": warnings, "failed_models": failed}


# ------------------------------------------------------------------ encoding/json
_HEX = "0123456789abcdef"


def go_json_string(s):
    """Go's encoding/json string encoder with HTML escaping (the default).  `s` is bytes or str;
    bytes may hold invalid UTF-8, each offending byte becomes \\ufffd."""
    b = s.encode("utf-8", "surrogatepass") if isinstance(s, str) else bytes(s)
    out = ['"']
    i, n = 0, len(b)
    while i < n:
        c = b[i]
        if c < 0x80:
            if c == 0x22:
                out.append('\\"')
            elif c == 0x5C:
                out.append("\\\\")
            elif c == 0x0A:
                out.append("\\n")
            elif c == 0x0D:
                out.append("\\r")
            elif c == 0x09:
                out.append("\\t")
            elif c == 0x08:
                out.append("\\b")          # go >= 1.22
            elif c == 0x0C:
                out.append("\\f")          # go >= 1.22
            elif c < 0x20 or c in (0x3C, 0x3E, 0x26):
                out.append("\\u00" + _HEX[c >> 4] + _HEX[c & 0xF])
            else:
                out.append(chr(c))
            i += 1
            continue
        # decode one UTF-8 sequence the way utf8.DecodeRune does
        size, cp = _decode_rune(b, i)
        if cp is None:
            out.append("\\ufffd")
            i += 1
            continue
        if cp in (0x2028, 0x2029):
            out.append("\\u202" + _HEX[cp & 0xF])
        else:
            out.append(b[i:i + size].decode("utf-8"))
        i += size
    out.append('"')
    return "".join(out)


def _decode_rune(b, i):
    c = b[i]
    n = len(b)
    if 0xC2 <= c <= 0xDF:
        need, lo, hi, cp = 1, 0x80, 0xBF, c & 0x1F
    elif c == 0xE0:
        need, lo, hi, cp = 2, 0xA0, 0xBF, c & 0x0F
    elif 0xE1 <= c <= 0xEC or 0xEE <= c <= 0xEF:
        need, lo, hi, cp = 2, 0x80, 0xBF, c & 0x0F
    elif c == 0xED:
        need, lo, hi, cp = 2, 0x80, 0x9F, c & 0x0F
    elif c == 0xF0:
        need, lo, hi, cp = 3, 0x90, 0xBF, c & 0x07
    elif 0xF1 <= c <= 0xF3:
        need, lo, hi, cp = 3, 0x80, 0xBF, c & 0x07
    elif c == 0xF4:
        need, lo, hi, cp = 3, 0x80, 0x8F, c & 0x07
    else:
        return 1, None
    for k in range(1, need + 1):
        if i + k >= n:
            return 1, None
        x = b[i + k]
        l, h = (lo, hi) if k == 1 else (0x80, 0xBF)
        if x < l or x > h:
            return 1, None
        cp = (cp << 6) | (x & 0x3F)
    return need + 1, cp


def encode_result(prompt, responses, consensus, judge, warnings=None, failed_models=None):
    """json.NewEncoder(w); SetIndent("", "  "); Encode(output.Result{...}).  responses: list of dicts with
    model, content, provider, latency_ms (int nanoseconds, the time.Duration value)."""
    js = go_json_string
    lines = ["{", '  "prompt": %s,' % js(prompt)]
    if responses is None:
        lines.append('  "responses": null,')
    elif not responses:
        lines.append('  "responses": [],')
    else:
        lines.append('  "responses": [')
        for k, r in enumerate(responses):
            lines.append("    {")
            lines.append('      "model": %s,' % js(r["model"]))
            lines.append('      "content": %s,' % js(r["content"]))
            lines.append('      "provider": %s,' % js(r["provider"]))
            lines.append('      "latency_ms": %d' % int(r["latency_ms"]))
            lines.append("    }" + ("," if k + 1 < len(responses) else ""))
        lines.append("  ],")
    lines.append('  "consensus": %s,' % js(consensus))
    tail = ['  "judge": %s' % js(judge)]
    for key, val in (("warnings", warnings), ("failed_models", failed_models)):
        if val:
            tail[-1] += ","
            tail.append('  "%s": [' % key)
            for k, w in enumerate(val):
                tail.append("    %s%s" % (js(w), "," if k + 1 < len(val) else ""))
            tail.append("  ]")
    lines.extend(tail)
    lines.append("}")
    return "\n".join(lines) + "\n"


# ------------------------------------------------------------------ token codec of the local provider
# Not part of the reference (it deals in strings only); restated here so that tests can check the
# C++ codec (llm-consensus_b200/host/codec.cpp) against an independent implementation.
def _plain(i):
    return i == 0x0A or (0x20 <= i <= 0x7E and i != 0x5B)      # '[' always escapes, so "[[" is unambiguous


def detokenize(ids):
    out = []
    for t in ids:
        out.append(chr(t) if _plain(t) else "[[%d]]" % t)
    return "".join(out)


def tokenize(text, vocab):
    b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
    ids, i, n = [], 0, len(b)
    while i < n:
        if b[i] == 0x5B and i + 1 < n and b[i + 1] == 0x5B:
            j = i + 2
            while j < n and j - (i + 2) < 7 and 0x30 <= b[j] <= 0x39:
                j += 1
            nd = j - (i + 2)
            if nd >= 1 and j + 1 < n and b[j] == 0x5D and b[j + 1] == 0x5D:
                digits = b[i + 2:j].decode()
                v = int(digits)
                if (nd == 1 or digits[0] != "0") and v < vocab and not _plain(v):
                    ids.append(v)
                    i = j + 2
                    continue
        ids.append(b[i])
        i += 1
    return ids
