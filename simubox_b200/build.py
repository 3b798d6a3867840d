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
    """Compile every source to an object file (in parallel: voronoi.cu alone takes ~50 s of template instances) and link
    them into libsimubox_b200.so.  Objects go to simubox_b200/build/ (git-ignored)."""
    if not force and not is_stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor

    obj_dir = PKG / "build"
    obj_dir.mkdir(exist_ok=True)
    nvcc = nvcc_path()
    common = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-O2,-ffp-contract=off", "-Xptxas", "-v" if verbose else "-O3", "-I", nccl_include()]
    srcs = [CSRC / s for s in SOURCES if (CSRC / s).exists()]

    def compile_one(src: Path):
        obj = obj_dir / (src.stem + ".o")
        cmd = [nvcc, "-c", str(src), "-o", str(obj), *common]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
        return obj, res.stdout + res.stderr

    with ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 4)) as pool:
        done = list(pool.map(compile_one, srcs))
    link = [nvcc, "-shared", "-o", str(LIB), *[str(o) for o, _ in done], "-gencode", "arch=compute_100a,code=sm_100a",
            "-Xcompiler", "-fPIC", "-lcufft", "-ldl"]
    res = subprocess.run(link, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + " ".join(link) + "\n" + res.stdout + res.stderr)
    if verbose:
        print("".join(log for _, log in done) + res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
