"""In-tree build of libbrp_b200.so (nvcc cross-compiles for sm_100a without a GPU).

    python boinc-app-eah-brp_b200/build.py            # or: __graft_entry__.build()
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "libbrp_b200.so"
CLI = HERE / "brp_b200_search"  # command-line driver over the batched API (csrc/brp_search_main.cpp)
SOURCES = ["resamp.cu", "fft.cu", "hs.cu", "search.cu", "legacy.cu", "noalloc.cu", "whiten.cu", "host.cpp", "zap.cpp"]
HEADERS = ["brp_internal.h", "brp_device.cuh", "ziggurat_tables.h", "../../include/brp_b200.h"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-fvisibility=default",
    "--expt-relaxed-constexpr",
]


PER_SOURCE_FLAGS: dict[str, list[str]] = {}
# tuning experiments: BRP_EXTRA_NVCC_FLAGS="-DBRP_HS_STAGES=3" python build.py --force
EXTRA_FLAGS = os.environ.get("BRP_EXTRA_NVCC_FLAGS", "").split()


def _stale(out: Path, deps: list[Path]) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = HERE / "build"
    objdir.mkdir(exist_ok=True)
    hdrs = [(CSRC / h).resolve() for h in HEADERS] + [Path(__file__)]
    objs, procs = [], []
    for src in SOURCES:
        obj = objdir / (src.rsplit(".", 1)[0] + ".o")
        objs.append(obj)
        if force or _stale(obj, [CSRC / src] + hdrs):
            cmd = [nvcc] + NVCC_FLAGS + EXTRA_FLAGS + PER_SOURCE_FLAGS.get(src, []) + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", str(CSRC / src), "-o", str(obj)]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", str(LIB)] + [str(o) for o in objs] + ["-lz", "-lcudart"]
        subprocess.run(cmd, check=True)
    main = CSRC / "brp_search_main.cpp"
    if force or _stale(CLI, [main, LIB, (CSRC / "../../include/brp_b200.h").resolve()]):
        cxx = os.environ.get("CXX", "g++")
        subprocess.run([cxx, "-O2", "-std=c++17", str(main), "-o", str(CLI), f"-L{HERE}", "-lbrp_b200",
                        "-Wl,-rpath,$ORIGIN", "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64", "-pthread"], check=True)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
