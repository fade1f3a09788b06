"""Build libllmc.so (the CUDA library + C ABI) in-tree for sm_100a.

    python llm-consensus_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The result (llm-consensus_b200/lib/libllmc.so) is
git-ignored but travels to the GPU box with the repository snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libllmc.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
         "-ccbin", "/usr/bin/g++", "--expt-relaxed-constexpr"]
# e.g. LLMC_NVCC_EXTRA="-DLLMC_RINGB_STATS" python llm-consensus_b200/build.py --force : the batched ring kernel with its
# cycle counters (profiles/ring_batch_trace.py prints them; they cost 4-5 % and are off in the normal build)
FLAGS += os.environ.get("LLMC_NVCC_EXTRA", "").split()


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _newest_header():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs += [os.path.join(HERE, "..", "include", f) for f in os.listdir(os.path.join(HERE, "..", "include"))]
    return max(os.path.getmtime(h) for h in hs)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    hdr = _newest_header()
    jobs = []
    objs = []
    for f in sources():
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f[:-3] + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr):
            cmd = [NVCC] + ARCH + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, r

    failed = False
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for cmd, r in ex.map(run, jobs):
            out = (r.stdout + r.stderr).strip()
            if r.returncode != 0:
                failed = True
                sys.stderr.write("FAILED: %s\n%s\n" % (" ".join(cmd), out))
            elif out and verbose:
                sys.stderr.write("%s\n%s\n" % (os.path.basename(cmd[-3]), out))
    if failed:
        raise RuntimeError("nvcc failed")
    if jobs or force or not os.path.exists(LIB):
        cmd = [NVCC] + ARCH + ["-shared", "-ccbin", "/usr/bin/g++", "-o", LIB] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    build_host(force or bool(jobs))
    return LIB


HOST = os.path.join(HERE, "host")
HOST_LIB = os.path.join(LIBDIR, "libllmc_host.so")
CLI = os.path.join(HERE, "bin", "llm-consensus")


def build_host(force=False):
    """The C++ host mirror: libllmc_host.so (test hooks) and the llm-consensus CLI (links libllmc.so)."""
    os.makedirs(os.path.dirname(CLI), exist_ok=True)
    srcs = [os.path.join(HOST, f) for f in ("host.cpp", "local_provider.cpp")]
    deps = srcs + [os.path.join(HOST, f) for f in ("host.hpp", "host_abi.cpp", "main.cpp")] + [LIB]
    newest = max(os.path.getmtime(d) for d in deps)
    cxx = ["/usr/bin/g++", "-O2", "-std=c++17", "-Wall", "-fPIC", "-pthread"]
    link = ["-L" + LIBDIR, "-lllmc", "-Wl,-rpath,$ORIGIN/../lib", "-Wl,-rpath,$ORIGIN"]
    if force or not os.path.exists(HOST_LIB) or os.path.getmtime(HOST_LIB) < newest:
        subprocess.check_call(cxx + ["-shared", "-o", HOST_LIB] + srcs + [os.path.join(HOST, "host_abi.cpp")] + link)
    if force or not os.path.exists(CLI) or os.path.getmtime(CLI) < newest:
        subprocess.check_call(cxx + ["-o", CLI] + srcs + [os.path.join(HOST, "main.cpp")] + link)
    return HOST_LIB, CLI


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
