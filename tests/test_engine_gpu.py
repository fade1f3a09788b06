"""End-to-end parity of the engine (prefill + greedy decode through the C ABI of include/llmc.h)
against the fp32 CPU oracle, on the same seeded random-init weights and the same token ids.

Tolerance (BASELINE.json north_star): logits max-abs <= 2e-2 against the fp32 oracle, teacher-forced on
the oracle's greedy tokens, and argmax agreement on the first 64 greedy tokens.  A free-running
disagreement is accepted ONLY where the oracle's own top-2 margin at that step is inside the logit
tolerance (a numerical tie that no bf16-activation implementation can be asked to resolve); the test
reports when that clause was needed."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# bring-up switch: LLMC_TEST_PREFILL_IMPL=1 runs every engine test on the plain CUDA-core GEMM
DEFAULT_IMPL = int(os.environ.get("LLMC_TEST_PREFILL_IMPL", "0"))


@pytest.fixture(scope="session")
def tc_ok(gpu):
    """The tcgen05 GEMM passes a small self-check in a subprocess (a hang there must not hang pytest)."""
    cmd = [sys.executable, os.path.join(ROOT, "tests", "_gemm_tc_check.py"), "130", "384", "256", "1", "0"]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    except subprocess.TimeoutExpired:
        return False
    if r.returncode != 0:
        return False
    return json.loads(r.stdout.strip().splitlines()[-1])["bad_rows"] == 0


@pytest.fixture(autouse=True)
def _guard_tensor_core_path(request, gpu):
    if DEFAULT_IMPL == 0 and not request.getfixturevalue("tc_ok"):
        pytest.fail("tcgen05 GEMM self-check failed or hung; engine tests on the product prefill path not run "
                    "(set LLMC_TEST_PREFILL_IMPL=1 to test the rest on the comparator GEMM)")

pytestmark = pytest.mark.gpu
TOL = 2e-2

DIMS = {k: v for k, v in ref.SHAPES.items() if k not in ("llama-1b", "llama-8b")}


def make_pair(llmc, name, seed, max_ctx=512, **kw):
    o = ref.Oracle(name, seed=seed, max_ctx=max_ctx)
    kw.setdefault("prefill_impl", DEFAULT_IMPL)
    e = llmc.Engine(dims=DIMS[name], seed=seed, max_ctx=max_ctx, max_slots=kw.pop("max_slots", 4), **kw)
    return o, e


@pytest.mark.parametrize("name", ["tiny", "tiny-untied", "small-d128"])
def test_weights_are_bit_identical_to_the_oracle(gpu, name):
    o, e = make_pair(gpu, name, seed=21)
    s = DIMS[name]
    for layer in [-1] + list(range(s["layers"])):
        kinds = [0, 1, 2] if layer < 0 else list(range(9))
        for kind in kinds:
            assert np.array_equal(e.debug_tensor(layer, kind), o.tensor(layer, kind)), (layer, kind)
    e.close()


@pytest.mark.parametrize("impl", [1, 0])
@pytest.mark.parametrize("name,n_prompt", [("tiny", 1), ("tiny", 37), ("tiny-untied", 70), ("small-d128", 45)])
def test_prefill_logits_and_kv(gpu, name, n_prompt, impl, request):
    if impl == 0 and not request.getfixturevalue("tc_ok"):
        pytest.fail("tcgen05 GEMM self-check failed or hung")
    o, e = make_pair(gpu, name, seed=3, prefill_impl=impl)
    V = DIMS[name]["vocab"]
    toks = np.random.default_rng(n_prompt).integers(0, V, n_prompt)
    want = o.forward(toks, "last")
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, toks, want_logits=True)
    err = np.abs(lg - want).max()
    assert err <= TOL, "prefill logits max-abs %.3e (std %.3f)" % (err, want.std())
    assert nxt == int(np.argmax(lg))
    assert e.seq_len(slot) == n_prompt
    e.seq_free(slot)
    e.close()


@pytest.mark.parametrize("flags", [0, 16, 19, 17, 18, 4, 20])   # 0: one-kernel decode; 16: 81 kernels per step (+1 no graph, +2 no PDL); 4: bf16 KV
@pytest.mark.parametrize("name", ["tiny", "tiny-untied", "small-d128"])
def test_teacher_forced_decode_logits(gpu, name, flags):
    """Every decode step's logits against the oracle, both fed the ORACLE's greedy tokens."""
    o, e = make_pair(gpu, name, seed=8, flags=flags)
    V = DIMS[name]["vocab"]
    prompt = np.random.default_rng(5).integers(0, V, 19)
    n_new = 40
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, prompt, want_logits=True)
    worst = np.abs(lg - olog[0]).max()
    agree = int(nxt == otoks[0])
    for s in range(1, n_new):
        e.set_next_token(slot, otoks[s - 1])
        out, lg = e.decode([slot], 1, want_logits=True)
        worst = max(worst, np.abs(lg[0] - olog[s]).max())
        assert out[0, 0] == int(np.argmax(lg[0]))
        agree += int(out[0, 0] == otoks[s])
    assert worst <= TOL, "decode logits max-abs %.3e" % worst
    assert e.seq_len(slot) == len(prompt) + n_new - 1
    # several steps in one call continue from the same device-resident state
    more = e.decode([slot], 5)
    assert more.shape == (1, 5) and e.seq_len(slot) == len(prompt) + n_new + 4
    print("teacher-forced argmax agreement %d/%d, worst logit err %.2e" % (agree, n_new, worst))
    e.close()


WIDE = ["wide-1b", "wide-d128", "wide-mha"]


@pytest.mark.parametrize("flags", [0, 128, 64, 256, 32, 4, 132, 68, 260, 36])   # 0: batched ring kernel with one sequence; 128: decode_ring (FMA slots, grid barriers); 64: tagged-word hand-overs; 256: mma.sync slots; 32: register-streaming kernel; 4: bf16 KV
@pytest.mark.parametrize("name", WIDE)
def test_ring_decode_teacher_forced_logits(gpu, name, flags):
    """The persistent decode kernels at hidden = 2048 (the ring kernel's shapes: rows of 2048 / 4096 / 8192
    elements, head sizes 64 / 128, GQA groups 4 / 1): every step's logits against the oracle, teacher-forced."""
    o, e = make_pair(gpu, name, seed=5, flags=flags, max_ctx=256, max_slots=1)
    V = DIMS[name]["vocab"]
    prompt = np.random.default_rng(11).integers(0, V, 45)
    n_new = 24
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, prompt, want_logits=True)
    worst = np.abs(lg - olog[0]).max()
    for s in range(1, n_new):
        e.set_next_token(slot, otoks[s - 1])
        out, lg = e.decode([slot], 1, want_logits=True)
        worst = max(worst, np.abs(lg[0] - olog[s]).max())
        assert out[0, 0] == int(np.argmax(lg[0]))
    tol = TOL if not flags & 4 else 6e-2    # bf16 KV is the reduced-precision mode (see test_reduced_precision_modes)
    assert worst <= tol, "decode logits max-abs %.3e" % worst
    print("%s flags %d: worst logit err %.2e" % (name, flags, worst))
    e.close()


@pytest.mark.parametrize("flags", [0, 128, 64, 256, 32])
def test_persistent_decode_long_context(gpu, flags):
    """Attention phase of the persistent kernels over a long context (2 300 positions: 72 tiles = the 8-way split cap,
    several tiles per warp, chunked prefill before it), teacher-forced against the oracle."""
    name = "wide-1b"
    o, e = make_pair(gpu, name, seed=6, flags=flags, max_ctx=2400, max_slots=1)
    V = DIMS[name]["vocab"]
    prompt = np.random.default_rng(21).integers(0, V, 2290)
    n_new = 10
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, prompt, want_logits=True)
    worst = np.abs(lg - olog[0]).max()
    for s in range(1, n_new):
        e.set_next_token(slot, otoks[s - 1])
        out, lg = e.decode([slot], 1, want_logits=True)
        worst = max(worst, np.abs(lg[0] - olog[s]).max())
    assert worst <= TOL, "logits max-abs %.3e" % worst
    e.close()


@pytest.mark.parametrize("name", WIDE)
@pytest.mark.parametrize("chunk", [1, 5, 64])
def test_ring_decode_free_running_matches_register_kernel(gpu, name, chunk):
    """Free-running greedy decode, many steps per call: ring kernel vs oracle, and token-identical to the
    register-streaming kernel (same arithmetic order per row is not guaranteed, so compare through the oracle's
    tie rule)."""
    o, e = make_pair(gpu, name, seed=14, max_ctx=256, max_slots=1)
    V = DIMS[name]["vocab"]
    prompt = np.random.default_rng(2).integers(0, V, 70)
    n_new = 64
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    gtoks = e.generate(prompt, n_new, chunk=chunk)
    where, margin = greedy_agreement(otoks, olog, gtoks)
    assert where < 0, "greedy tokens diverge at step %d with oracle margin %.3e > tol" % (where, margin)
    e.close()


TIES = {"count": 0}     # free-running divergences at oracle near-ties seen in this session (reported, never silent)


def greedy_agreement(otoks, olog, gtoks, tol=TOL):
    """Index of the first free-running disagreement that is NOT a numerical tie in the oracle, or -1."""
    for s, (a, b) in enumerate(zip(otoks, gtoks)):
        if a != b:
            margin = float(olog[s][a] - olog[s][b])
            if margin > tol:
                return (s, margin)
            TIES["count"] += 1
            print("NOTE: free-running divergence at step %d on an oracle near-tie (top-2 margin %.2e <= %.0e); "
                  "%d such case(s) so far in this session" % (s, margin, tol, TIES["count"]))
            return (-2, margin)                                        # -2: diverged at a tie
    return (-1, 0.0)


def check_batched(o, prompts, firsts, out, lg, n_new, tol=TOL, which=None):
    """Every sequence of a batched decode call against its own oracle run: greedy tokens (ties excepted, and
    reported), and the last step's logits.  A sequence that left the oracle's path at a tie is NOT dropped: the
    oracle is re-run on the GPU's own tokens (teacher forcing) and the last-step logits are compared on that path."""
    worst, ties = 0.0, 0
    for i in (which if which is not None else range(len(prompts))):
        o.reset()
        otoks, olog = o.generate(prompts[i], n_new + 1, want_logits=True)
        gt = [firsts[i]] + out[i].tolist()
        where, margin = greedy_agreement(otoks, olog, gt, tol=tol)
        assert where < 0, (i, where, margin)
        if where == -1:
            want = olog[n_new]
        else:
            ties += 1
            o.reset()
            want = o.forward(np.concatenate([np.asarray(prompts[i]), np.asarray(gt[:n_new])]), "last")
        err = float(np.abs(lg[i] - want).max())
        assert err <= tol, "sequence %d: last-step logits max-abs %.3e (tie path: %s)" % (i, err, where == -2)
        worst = max(worst, err)
    return worst, ties


@pytest.mark.parametrize("name", ["tiny", "tiny-untied", "small-d128"])
@pytest.mark.parametrize("chunk,flags", [(1, 0), (7, 0), (64, 0), (7, 16), (64, 16)])
def test_free_running_greedy_tokens(gpu, name, chunk, flags):
    o, e = make_pair(gpu, name, seed=13, flags=flags)
    V = DIMS[name]["vocab"]
    prompt = np.random.default_rng(9).integers(0, V, 33)
    n_new = 64
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    gtoks = e.generate(prompt, n_new, chunk=chunk)
    where, margin = greedy_agreement(otoks, olog, gtoks)
    assert where < 0, "greedy tokens diverge at step %d with oracle margin %.3e > tol" % (where, margin)
    if where == -2:
        print("NOTE: free-running divergence at an oracle near-tie (margin %.2e <= %.0e)" % (margin, TOL))
    e.close()


@pytest.mark.parametrize("name", ["tiny", "small-d128"])
@pytest.mark.parametrize("flags", [4, 8, 12])
def test_reduced_precision_modes(gpu, name, flags):
    """LLMC_FLAG_KV_BF16 (4) / LLMC_FLAG_PREFILL_BF16 (8): bf16 storage of the KV cache and of the prefill
    GEMM operands.  On these small shapes the result stays inside the 2e-2 budget; for the KV flag the
    deviation from the fp32 oracle must moreover be explained by exactly that rounding (the oracle's
    diagnostic emulation of bf16 KV storage reproduces the GPU logits ~20x more closely than the fp32
    oracle does); for the prefill flag the emulation is only required not to be worse."""
    o, e = make_pair(gpu, name, seed=17, flags=flags)
    V = DIMS[name]["vocab"]
    prompt = np.random.default_rng(3).integers(0, V, 60)
    mask = (4 if flags & 4 else 0) | ((1 | 8 | 16) if flags & 8 else 0)
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, prompt, want_logits=True)
    exact = o.forward(prompt, "last")
    o.reset(); o.set_round_mask(mask)
    emu = o.forward(prompt, "last")
    err_exact, err_emu = np.abs(lg - exact).max(), np.abs(lg - emu).max()
    print("flags %d: vs fp32 oracle %.2e, vs storage-rounding emulation %.2e" % (flags, err_exact, err_emu))
    assert err_exact <= TOL                      # small shapes stay inside the budget even so
    assert err_emu <= (1e-3 if flags == 4 else max(2e-3, err_exact))
    # a few decode steps, teacher-forced on the emulation's tokens
    for s_ in range(6):
        tok = int(np.argmax(emu))
        e.set_next_token(slot, tok)
        _, lgd = e.decode([slot], 1, want_logits=True)
        emu = o.forward([tok], "last")
        assert np.abs(lgd[0] - emu).max() <= (1e-3 if flags == 4 else TOL)
    e.close()


def test_chunked_prefill_equals_single_pass(gpu):
    o, e = make_pair(gpu, "tiny", seed=2, prefill_chunk=32)
    _, e2 = make_pair(gpu, "tiny", seed=2, prefill_chunk=2048)
    toks = np.random.default_rng(1).integers(0, 1024, 150)
    s1, s2 = e.seq_alloc(), e2.seq_alloc()
    n1, l1 = e.prefill(s1, toks, want_logits=True)        # 5 chunks of 32 (ragged last)
    n2, l2 = e2.prefill(s2, toks, want_logits=True)
    want = o.forward(toks, "last")
    assert np.abs(l1 - want).max() <= TOL and np.abs(l2 - want).max() <= TOL
    assert np.abs(l1 - l2).max() <= 1e-3
    # split prompt over two llmc_prefill calls (prefix reuse)
    s3 = e.seq_alloc()
    e.prefill(s3, toks[:70])
    n3, l3 = e.prefill(s3, toks[70:], want_logits=True)
    assert np.abs(l3 - want).max() <= TOL and e.seq_len(s3) == 150
    e.close(); e2.close()


@pytest.mark.parametrize("B", [2, 3, 5, 8, 11, 16, 33, 128, 131])   # <= 8: GEMV kernels; 9..128: wide (GEMM) path; > 128: groups
def test_batched_decode_matches_per_sequence_oracle(gpu, B):
    name = "tiny"
    V = DIMS[name]["vocab"]
    e = gpu.Engine(dims=DIMS[name], seed=4, max_ctx=256, max_slots=max(12, B), prefill_impl=DEFAULT_IMPL)
    rng = np.random.default_rng(B)
    prompts = [rng.integers(0, V, int(n)) for n in rng.integers(3, 90, B)]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    n_new = 12
    out, lg = e.decode(slots, n_new, want_logits=True)
    o = ref.Oracle(name, seed=4, max_ctx=256)
    worst, ties = check_batched(o, prompts, firsts, out, lg, n_new)
    print("batch %d: worst last-step logit err %.2e, %d sequence(s) checked on their own path after a tie" % (B, worst, ties))
    e.close()


def test_slots_are_independent_and_reusable(gpu):
    e = gpu.Engine(dims=DIMS["tiny"], seed=6, max_ctx=128, max_slots=2, prefill_impl=DEFAULT_IMPL)
    p = np.arange(1, 30)
    a = e.generate(p, 10)
    s0 = e.seq_alloc(); s1 = e.seq_alloc()
    with pytest.raises(gpu.LlmcError) as ei:
        e.seq_alloc()
    assert ei.value.code == gpu.ERR_LIMIT
    e.prefill(s0, [5, 6, 7])
    e.seq_free(s0); e.seq_free(s1)
    b = e.generate(p, 10)                                  # same pages, recycled
    assert a.tolist() == b.tolist()
    e.close()


@pytest.mark.parametrize("flags", [0, 4])   # fp32 and bf16 KV pages
def test_seq_fork_clone_prefix_and_page_sharing(gpu, flags):
    """llmc_seq_fork: (1) a clone at the source's end decodes bit-identically to the source; (2) a fork of a prefix,
    fed the rest of the prompt, matches the single-pass prefill; (3) shared full pages survive freeing either side and
    the copied tail page keeps the two sequences' appended rows apart; (4) nothing leaks from the page pool."""
    name = "tiny"
    V = DIMS[name]["vocab"]
    e = gpu.Engine(dims=DIMS[name], seed=6, max_ctx=256, max_slots=4, prefill_impl=DEFAULT_IMPL, flags=flags)
    rng = np.random.default_rng(17)
    prompt = rng.integers(0, V, 150)            # 2 full pages of 64 + 22 positions in the third
    a = e.seq_alloc()
    first, lg_full = e.prefill(a, prompt, want_logits=True)
    b = e.seq_fork(a)                           # clone: 2 shared pages, copied tail, pending token inherited
    assert e.seq_len(b) == 150 and e.get_next_token(b) == first
    c = e.seq_fork(a, 100)                      # prefix: 1 shared page + 36 copied positions
    assert e.seq_len(c) == 100
    first_c, lg_c = e.prefill(c, prompt[100:], want_logits=True)
    tol = TOL if not flags & 12 else 8e-2
    assert np.abs(lg_c - lg_full).max() <= tol and e.seq_len(c) == 150
    # the source and the clone decode one after the other (each appends to its own copy of the tail page)
    out_a, lg_a = e.decode([a], 12, want_logits=True)
    out_b, lg_b = e.decode([b], 12, want_logits=True)
    assert out_a.tolist() == out_b.tolist() and np.array_equal(lg_a, lg_b)
    # the source goes away; the clone still reads the pages they shared
    e.seq_free(a)
    d = e.seq_alloc()
    assert e.prefill(d, prompt) == first
    out_d1 = e.decode([d], 12)
    out_b2, lg_b2 = e.decode([b], 6, want_logits=True)
    out_d2, lg_d2 = e.decode([d], 6, want_logits=True)
    assert out_d1.tolist() == out_a.tolist() and out_b2.tolist() == out_d2.tolist() and np.array_equal(lg_b2, lg_d2)
    # oracle check of the clone's path
    o = ref.Oracle(name, seed=6, max_ctx=256)
    otoks, olog = o.generate(prompt, 19, want_logits=True)
    gt = [first] + out_b[0].tolist() + out_b2[0].tolist()
    where, margin = greedy_agreement(otoks, olog, gt, tol=tol)
    assert where < 0, (where, margin)
    # errors, and the pool: every page comes back
    for bad in (0, 151 + 18):
        with pytest.raises(gpu.LlmcError):
            e.seq_fork(b, bad)
    with pytest.raises(gpu.LlmcError):          # all four slots are in use: b, c, d and this one
        x = e.seq_alloc()
        e.seq_fork(b)
    for s_ in (b, c, d, x):
        e.seq_free(s_)
    for rep in range(3):                        # 4 slots x 4 pages = the whole pool, three times over
        ss = [e.seq_alloc() for _ in range(2)]
        for s_ in ss:
            e.prefill(s_, rng.integers(0, V, 250))
        ff = [e.seq_fork(s_, 200) for s_ in ss]
        for s_ in ff:
            e.prefill(s_, rng.integers(0, V, 50))
        for s_ in ss + ff:
            e.seq_free(s_)
    e.close()


def test_error_paths(gpu):
    e = gpu.Engine(dims=DIMS["tiny"], seed=6, max_ctx=64, max_slots=2, prefill_impl=DEFAULT_IMPL)
    s = e.seq_alloc()
    with pytest.raises(gpu.LlmcError) as ei:
        e.decode([s], 1)                                   # nothing prefilled
    assert ei.value.code == gpu.ERR_STATE
    with pytest.raises(gpu.LlmcError) as ei:
        e.prefill(s, [5000])                               # token out of range
    assert ei.value.code == gpu.ERR_INVALID
    with pytest.raises(gpu.LlmcError) as ei:
        e.prefill(s, list(range(65)))                      # beyond max_ctx
    assert ei.value.code == gpu.ERR_LIMIT
    e.prefill(s, list(range(60)))
    with pytest.raises(gpu.LlmcError) as ei:
        e.decode([s], 10)                                  # would overflow the context
    assert ei.value.code == gpu.ERR_LIMIT
    with pytest.raises(gpu.LlmcError) as ei:
        e.decode([s, s], 1)
    assert ei.value.code == gpu.ERR_INVALID
    with pytest.raises(gpu.LlmcError) as ei:
        e.seq_free(1)
    assert ei.value.code == gpu.ERR_STATE
    assert e.decode([s], 4).shape == (1, 4)                # still usable after the errors
    e.close()


def test_llama8b_shape_short(gpu):
    """The 8B shape (head_dim 128, K 4096 / 14336, untied LM head) at full width on a short sequence:
    teacher-forced logits against the oracle, through both decode paths."""
    o = ref.Oracle("llama-8b", seed=2, max_ctx=64)
    prompt = np.random.default_rng(3).integers(0, 128256, 40)
    n_new = 4
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    for flags in (0, 16):
        e = gpu.Engine(shape=gpu.SHAPE_LLAMA_8B, seed=2, max_ctx=64, max_slots=1, prefill_impl=DEFAULT_IMPL, flags=flags)
        slot = e.seq_alloc()
        nxt, lg = e.prefill(slot, prompt, want_logits=True)
        worst = float(np.abs(lg - olog[0]).max())
        for s in range(1, n_new):
            e.set_next_token(slot, otoks[s - 1])
            out, lg = e.decode([slot], 1, want_logits=True)
            worst = max(worst, float(np.abs(lg[0] - olog[s]).max()))
        print("8B shape flags %d: worst logit err %.2e (logit std %.3f)" % (flags, worst, float(olog.std())))
        assert worst <= TOL
        e.close()


def test_config2_llama1b_512_prompt_256_decode(gpu):
    """BASELINE.json configs[1] at full size: Llama-3.2-1B shape, seed 0, 512-token prompt, 256 greedy
    tokens; logits teacher-forced on the oracle's tokens, every step, max-abs <= 2e-2; argmax
    agreement on the first 64 free-running tokens (numerical ties in the oracle excepted)."""
    n_prompt, n_new = 512, 256
    o = ref.Oracle("llama-1b", seed=0, max_ctx=1024)
    prompt = np.random.default_rng(1).integers(0, 128256, n_prompt)
    otoks, olog = o.generate(prompt, n_new, want_logits=True)
    e = gpu.Engine(shape=gpu.SHAPE_LLAMA_1B, seed=0, max_ctx=1024, max_slots=2, prefill_impl=DEFAULT_IMPL)
    # free-running
    gtoks = e.generate(prompt, n_new)
    where, margin = greedy_agreement(otoks[:64], olog[:64], gtoks[:64])
    n_same = int((np.asarray(gtoks) == otoks).sum())
    print("free-running: %d/%d tokens equal; first-64 check -> %s (margin %.3e)" % (n_same, n_new, where, margin))
    assert where < 0, "first 64 greedy tokens diverge at %d with oracle margin %.3e" % (where, margin)
    # teacher-forced logits
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, prompt, want_logits=True)
    errs = [float(np.abs(lg - olog[0]).max())]
    agree = int(nxt == otoks[0])
    for s in range(1, n_new):
        e.set_next_token(slot, otoks[s - 1])
        out, lg = e.decode([slot], 1, want_logits=True)
        errs.append(float(np.abs(lg[0] - olog[s]).max()))
        agree += int(out[0, 0] == otoks[s])
    print("teacher-forced: worst logit err %.3e (prefill %.3e), argmax agreement %d/%d, logit std %.3f"
          % (max(errs), errs[0], agree, n_new, float(olog.std())))
    assert max(errs) <= TOL
    st = e.stats()
    assert st["kernel_launches"] > 0
    e.close()


def test_wide_batch_decode_full_size_1b(gpu):
    """The wide decode path (tcgen05 skinny GEMMs, split-K, batched RoPE/append, two-stage argmax) at the full
    Llama-3.2-1B shape: 12 sequences of different lengths decoded together, each against its own oracle run."""
    B, n_new = 12, 6
    o = ref.Oracle("llama-1b", seed=0, max_ctx=128)
    e = gpu.Engine(shape=gpu.SHAPE_LLAMA_1B, seed=0, max_ctx=128, max_slots=B, prefill_impl=DEFAULT_IMPL)
    rng = np.random.default_rng(77)
    prompts = [rng.integers(0, 128256, int(n)) for n in rng.integers(5, 60, B)]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    out, lg = e.decode(slots, n_new, want_logits=True)
    worst, ties = check_batched(o, prompts, firsts, out, lg, n_new)
    print("wide decode, 1B shape, %d sequences: worst last-step logit err %.3e (%d on their own path after a tie)" % (B, worst, ties))
    assert worst <= TOL
    e.close()


GOLDEN = os.path.join(ROOT, "tests", "golden")


def fixture_err(lg_row, probe_idx, probe_val, top_idx, top_val):
    """max-abs difference of one GPU logit row against the oracle values a committed fixture holds for it: the logits
    at 512 fixed vocabulary positions and the oracle's top 16."""
    return max(float(np.abs(lg_row[probe_idx] - probe_val).max()), float(np.abs(lg_row[top_idx] - top_val).max()))


@pytest.mark.parametrize("B", [32, 128])
def test_wide_batch_decode_full_size_1b_b32_b128(gpu, B):
    """BASELINE.json configs[4]'s 1B points at B = 32 and 128 (contexts 512-560): the tcgen05 decode GEMMs of the wide
    path -- gemm_decode_tc.cu up to 32 sequences, gemm_tcgen05.cu in decode mode with split-K above -- at the full
    Llama-3.2-1B shape.  Three sequences of each batch are checked against their own fp32 oracle run, replayed from
    tests/golden/oracle_1b_wide.npz (generator: tests/golden/gen_oracle_1b_wide_fixture.py, which also defines the
    prompts)."""
    path = os.path.join(GOLDEN, "oracle_1b_wide.npz")
    if not os.path.exists(path):
        pytest.fail("fixture %s is missing; run tests/golden/gen_oracle_1b_wide_fixture.py" % path)
    sys.path.insert(0, GOLDEN)
    import gen_oracle_1b_wide_fixture as gen
    fx = np.load(path)
    n_new = int(fx["n_new"])
    prompts = gen.prompts_for(B)
    e = gpu.Engine(shape=gpu.SHAPE_LLAMA_1B, seed=0, max_ctx=640, max_slots=B, prefill_impl=DEFAULT_IMPL)
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    out, lg = e.decode(slots, n_new, want_logits=True)
    worst, ties = 0.0, 0
    for i in gen.CHECK[B]:
        otoks, margins = fx["b%d_s%d_tokens" % (B, i)], fx["b%d_s%d_margins" % (B, i)]
        gt = [firsts[i]] + out[i].tolist()
        same_path = True
        for s_, (a, b) in enumerate(zip(otoks.tolist(), gt)):
            if a != b:
                assert margins[s_] <= TOL, "sequence %d diverges at step %d where the oracle margin is %.3e" % (i, s_, margins[s_])
                print("NOTE: B=%d sequence %d left the oracle's path at step %d on a near-tie (margin %.2e)" % (B, i, s_, margins[s_]))
                same_path, ties = False, ties + 1
                break
        if same_path:
            err = fixture_err(lg[i], fx["probe_idx"], fx["b%d_s%d_probe" % (B, i)], fx["b%d_s%d_top_idx" % (B, i)],
                              fx["b%d_s%d_top_val" % (B, i)])
            assert err <= TOL, "sequence %d: last-step logits differ by %.3e" % (i, err)
            worst = max(worst, err)
    print("wide decode, 1B shape, B=%d, contexts 512-566: worst checked logit err %.3e, %d tie(s)" % (B, worst, ties))
    assert ties < len(gen.CHECK[B])
    e.close()


def test_llama8b_2048_prompt_fixture(gpu):
    """BASELINE.json configs[3]'s model at its prompt length: the Llama-3-8B shape, 2 048-token prompt (one full
    prefill chunk, 32 key tiles per head in the tensor-core prefill attention), then 64 decode steps teacher-forced on
    the oracle's tokens (contexts 2 049..2 112: the decode attention runs split over several blocks).  The fp32 oracle
    run (tens of CPU-minutes) is replayed from tests/golden/oracle_8b_2048.npz (tests/golden/gen_oracle_8b_fixture.py)."""
    path = os.path.join(GOLDEN, "oracle_8b_2048.npz")
    if not os.path.exists(path):
        pytest.fail("fixture %s is missing; run tests/golden/gen_oracle_8b_fixture.py" % path)
    fx = np.load(path)
    n_prompt, otoks = int(fx["n_prompt"]), fx["tokens"]
    prompt = np.random.default_rng(int(fx["prompt_seed"])).integers(0, 128256, n_prompt).astype(np.int32)
    e = gpu.Engine(shape=gpu.SHAPE_LLAMA_8B, seed=0, max_ctx=n_prompt + len(otoks) + 8, max_slots=1, prefill_impl=DEFAULT_IMPL)
    slot = e.seq_alloc()
    nxt, lg = e.prefill(slot, prompt, want_logits=True)
    errs = [fixture_err(lg, fx["probe_idx"], fx["probe_val"][0], fx["top_idx"][0], fx["top_val"][0])]
    margin = fx["top_val"][:, 0] - fx["top_val"][:, 1]
    agree = int(nxt == otoks[0])
    assert nxt == otoks[0] or margin[0] <= TOL
    for s_ in range(1, len(otoks)):
        e.set_next_token(slot, int(otoks[s_ - 1]))
        out, lg = e.decode([slot], 1, want_logits=True)
        errs.append(fixture_err(lg[0], fx["probe_idx"], fx["probe_val"][s_], fx["top_idx"][s_], fx["top_val"][s_]))
        agree += int(out[0, 0] == otoks[s_])
        assert out[0, 0] == otoks[s_] or margin[s_] <= TOL, "step %d: argmax differs where the oracle margin is %.3e" % (s_, margin[s_])
    print("8B shape, 2048-token prompt + %d teacher-forced steps: worst logit err %.3e (prefill %.3e), argmax agreement %d/%d, "
          "logit std %.3f" % (len(otoks) - 1, max(errs), errs[0], agree, len(otoks), float(fx["logit_std"])))
    assert max(errs) <= TOL
    e.close()


@pytest.mark.parametrize("flags", [1, 2, 3, 4, 8, 12])   # no graph, no PDL, both, bf16 KV, single-bf16 operands, both
@pytest.mark.parametrize("B", [5, 40])
def test_wide_decode_engine_flags(gpu, B, flags):
    """The wide decode path under every engine flag that changes how its kernels are launched or what they read."""
    name = "small-d128"
    V = DIMS[name]["vocab"]
    e = gpu.Engine(dims=DIMS[name], seed=9, max_ctx=128, max_slots=B, prefill_impl=DEFAULT_IMPL, flags=flags)
    rng = np.random.default_rng(B + flags)
    prompts = [rng.integers(0, V, int(n)) for n in rng.integers(3, 50, B)]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    n_new = 8
    out, lg = e.decode(slots, n_new, want_logits=True)
    o = ref.Oracle(name, seed=9, max_ctx=128)
    tol = TOL if not flags & 12 else 8e-2      # reduced-precision storage modes: see test_reduced_precision_modes
    check_batched(o, prompts, firsts, out, lg, n_new, tol=tol, which=range(0, B, max(1, B // 6)))
    e.close()


@pytest.mark.parametrize("flags", [0, 4, 128])     # fp32 KV, bf16 KV, and the per-step kernel graph as the comparator
@pytest.mark.parametrize("name,B", [("wide-1b", 2), ("wide-1b", 3), ("wide-1b", 4), ("wide-d128", 2), ("wide-d128", 4)])
def test_ring_batch_decode(gpu, name, B, flags):
    """2..4 sequences through the persistent batched ring kernel (decode_ring_batch.cu): sequences of different
    lengths (one context long enough for several attention splits), two decode calls in a row (the second starts from
    the state the first left on the device), every sequence against its own oracle run."""
    V = DIMS[name]["vocab"]
    e = gpu.Engine(dims=DIMS[name], seed=21, max_ctx=512, max_slots=4, prefill_impl=DEFAULT_IMPL, flags=flags)
    o = ref.Oracle(name, seed=21, max_ctx=512)
    rng = np.random.default_rng(B * 7 + flags)
    lens = [330, 5, 97, 140][:B]
    prompts = [rng.integers(0, V, n) for n in lens]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    out1 = e.decode(slots, 5)
    out2, lg = e.decode(slots, 4, want_logits=True)
    out = [np.concatenate([out1[i], out2[i]]) for i in range(B)]
    tol = TOL if not flags & 12 else 8e-2
    worst, ties = check_batched(o, prompts, firsts, out, lg, 9, tol=tol)
    assert [e.seq_len(s) for s in slots] == [n + 9 for n in lens]
    print("%s, %d sequences, flags %d: worst last-step logit err %.3e (%d on their own path after a tie)" % (name, B, flags, worst, ties))
    e.close()


def test_ring_batch_decode_long_and_uneven_contexts(gpu):
    """Three sequences through the batched ring kernel with very different contexts -- 2 250, 33 and 1 100 positions:
    the attention splits per (sequence, kv head) differ per sequence (6 blocks each are available at batch 3), the
    short one runs unsplit -- over a chunk of 7 steps, each against its own oracle run."""
    name = "wide-1b"
    V = DIMS[name]["vocab"]
    e = gpu.Engine(dims=DIMS[name], seed=23, max_ctx=2304, max_slots=3, prefill_impl=DEFAULT_IMPL)
    o = ref.Oracle(name, seed=23, max_ctx=2304)
    rng = np.random.default_rng(31)
    lens = [2250, 33, 1100]
    prompts = [rng.integers(0, V, n) for n in lens]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    out, lg = e.decode(slots, 7, want_logits=True)
    worst, ties = check_batched(o, prompts, firsts, out, lg, 7)
    print("uneven contexts %s: worst last-step logit err %.3e (%d on their own path after a tie)" % (lens, worst, ties))
    e.close()


def test_ring_batch_decode_full_size_1b(gpu):
    """The batched ring kernel at the full Llama-3.2-1B shape: 4 sequences, a long and three short contexts, against the
    oracle; then the same slots continue as two pairs (batch 2) and must produce the tokens the oracle continues with."""
    n_new = 6
    o = ref.Oracle("llama-1b", seed=0, max_ctx=256)
    e = gpu.Engine(shape=gpu.SHAPE_LLAMA_1B, seed=0, max_ctx=256, max_slots=4, prefill_impl=DEFAULT_IMPL)
    rng = np.random.default_rng(78)
    prompts = [rng.integers(0, 128256, n) for n in (150, 7, 40, 66)]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    out, lg = e.decode(slots, n_new, want_logits=True)
    worst, ties = check_batched(o, prompts, firsts, out, lg, n_new)
    print("ring batch, 1B shape, 4 sequences: worst last-step logit err %.3e (%d on their own path after a tie)" % (worst, ties))
    assert worst <= TOL
    e.close()


def test_wide_decode_long_context_two_attention_blocks(gpu):
    """Wide path with contexts past 2 048 tokens: the decode attention then runs two blocks per (sequence, kv head);
    the one that owns the last tile applies RoPE to the new K row and appends K/V itself."""
    name = "tiny"
    V = DIMS[name]["vocab"]
    e = gpu.Engine(dims=DIMS[name], seed=12, max_ctx=2304, max_slots=3, prefill_impl=DEFAULT_IMPL)
    o = ref.Oracle(name, seed=12, max_ctx=2304)
    rng = np.random.default_rng(5)
    prompts = [rng.integers(0, V, n) for n in (2100, 2047, 2200)]
    slots, firsts = [], []
    for p in prompts:
        s = e.seq_alloc()
        slots.append(s)
        firsts.append(e.prefill(s, p))
    n_new = 6
    out, lg = e.decode(slots, n_new, want_logits=True)
    check_batched(o, prompts, firsts, out, lg, n_new)
    e.close()
