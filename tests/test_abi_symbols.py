"""The C-ABI library loads on a CPU-only machine and exports every symbol include/*.h declares;
without a GPU it refuses to create an engine (there is no CPU fallback)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(llmc_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(llmc):
    names = declared("llmc.h") + declared("llmc_ops.h")
    assert len(names) >= 20
    for n in names:
        assert hasattr(llmc.lib(), n), "libllmc.so does not export " + n
    assert sorted(llmc.ENGINE_SYMBOLS) == declared("llmc.h")
    assert sorted(llmc.OPS_SYMBOLS) == declared("llmc_ops.h")


def test_host_library_exports_its_header(llmc):
    import ctypes
    import llm_consensus_b200.build as b
    lib_path, cli = b.build_host()
    h = ctypes.CDLL(lib_path)
    names = declared("llmc_host.h")
    assert len(names) == 5
    for n in names:
        assert hasattr(h, n), "libllmc_host.so does not export " + n
    assert os.access(cli, os.X_OK)


def test_no_torch_or_oracle_in_the_product_library(llmc):
    import subprocess
    out = subprocess.run(["ldd", llmc.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "oracle" not in out and "nccl" not in out
    # product sources never reference the oracle
    pkg = os.path.join(ROOT, "llm-consensus_b200")
    for dp, _, fs in os.walk(pkg):
        if os.sep + "build" in dp:
            continue
        for f in fs:
            if f.endswith((".cu", ".cuh", ".h", ".cpp", ".hpp", ".py")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_create_fails_loudly_without_gpu(llmc):
    if llmc.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(llmc.LlmcError) as ei:
        llmc.Engine(dims=dict(hidden=256, layers=1, q_heads=8, kv_heads=2, head_dim=32, inter=512, vocab=1024, tied=1))
    assert ei.value.code == llmc.ERR_CUDA and "no CPU fallback" in str(ei.value)


def test_config_validation_happens_before_device_use(llmc):
    bad = dict(hidden=250, layers=1, q_heads=8, kv_heads=2, head_dim=32, inter=512, vocab=1024, tied=1)
    with pytest.raises(llmc.LlmcError) as ei:
        llmc.Engine(dims=bad)
    assert ei.value.code == llmc.ERR_INVALID
    with pytest.raises(llmc.LlmcError) as ei:
        llmc.Engine(shape=77)
    assert ei.value.code == llmc.ERR_INVALID


def test_go_binding_source_matches_the_document_and_the_abi(llmc):
    """The cgo binding kept under llm-consensus_b200/go/ (uncompiled: no Go toolchain here) is the block shown in
    INTEGRATION.md, and every C name it calls is a symbol libllmc.so exports."""
    god = os.path.join(ROOT, "llm-consensus_b200", "go", "internal", "provider")
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    a = doc.index("```go\n//go:build cuda") + len("```go\n")
    block = doc[a:doc.index("```", a)]
    src = open(os.path.join(god, "local.go")).read()
    assert src == block
    called = sorted(set(re.findall(r"\bC\.(llmc_[a-z0-9_]+)\(", src)))
    assert "llmc_engine_create" in called and "llmc_decode" in called and len(called) >= 7
    for n in called:
        assert hasattr(llmc.lib(), n), "binding calls C.%s, which libllmc.so does not export" % n
    for f in ("local_stub.go", "local_codec.go"):
        assert os.path.exists(os.path.join(god, f))


def test_go_codec_rules_match_the_host_codec():
    """local_codec.go restates host.cpp's codec; a Python transcription of the Go text's rules is compared with the
    C++ implementation (through libllmc_host.so) on awkward strings."""
    import ctypes
    import llm_consensus_b200.build as b
    lib_path, _ = b.build_host()
    h = ctypes.CDLL(lib_path)

    def plain(t):
        return t == 0x0A or (0x20 <= t <= 0x7E and t != ord("["))

    def tokenize(text, vocab):
        bts, ids, i, n = text.encode("latin-1"), [], 0, len(text)
        while i < n:
            if bts[i] == 91 and i + 1 < n and bts[i + 1] == 91:
                j = i + 2
                while j < n and j - (i + 2) < 7 and 48 <= bts[j] <= 57:
                    j += 1
                nd = j - (i + 2)
                if nd >= 1 and j + 1 < n and bts[j] == 93 and bts[j + 1] == 93:
                    v = int(bts[i + 2:j])
                    if (nd == 1 or bts[i + 2] != 48) and v < vocab and not plain(v):
                        ids.append(v)
                        i = j + 2
                        continue
            ids.append(bts[i])
            i += 1
        return ids

    h.llmc_host_tokenize.restype = ctypes.c_int64
    for text in ["plain text\n", "[[5]] [[05]] [[65]] [[128255]] [[128256]] [[", "a[[12", "[[1234567]][[12345678]]", "\x00\x7f\xff[]]"]:
        raw = text.encode("latin-1")
        cap = len(raw) + 8
        out = (ctypes.c_int32 * cap)()
        n = h.llmc_host_tokenize(raw, len(raw), 128256, out, cap)
        if n < 0:
            pytest.skip("host tokenize signature differs")
        assert list(out[:n]) == tokenize(text, 128256), text
