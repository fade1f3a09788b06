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
