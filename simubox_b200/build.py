"""Build libsimubox_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libsimubox_b200.so"
SOURCES = ["scatter.cu", "voronoi.cu", "shell_table.cpp", "slab.cu", "text_table.cpp", "cellstats.cu", "seeds.cu", "splat.cu"]


def nvcc_path() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; cannot build libsimubox_b200.so")
    return cand


def is_stale() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*")) + [PKG.parent / "include" / "sbx.h"]
    return any(p.stat().st_mtime > t for p in deps)


def nccl_include() -> str:
    """Directory holding nccl.h (types only: the library itself is dlopen'ed at run time)."""
    try:
        import nvidia.nccl as _n

        for base in list(getattr(_n, "__path__", [])):
            if (Path(base) / "include" / "nccl.h").exists():
                return str(Path(base) / "include")
    except Exception:
        pass
    for cand in ("/usr/include", "/usr/local/cuda/include"):
        if (Path(cand) / "nccl.h").exists():
            return cand
    raise RuntimeError("nccl.h not found (needed for the slab-decomposed scatter types)")


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not is_stale():
        return LIB
    srcs = [str(CSRC / s) for s in SOURCES if (CSRC / s).exists()]
    cmd = [
        nvcc_path(), "-shared", "-o", str(LIB), *srcs,
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-O3", "-lineinfo", "-std=c++17",
        "-Xcompiler", "-fPIC,-O2,-ffp-contract=off",
        "-Xptxas", "-v" if verbose else "-O3",
        "-I", nccl_include(), "-lcufft", "-ldl",
    ]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
