"""In-tree build of libxgnn_b200.so (sm_100a only) and of the test-side oracle libraries.

`python -m xgnn_b200.build` or `xgnn_b200.build.build_all()`.  nvcc cross-compiles without a GPU.
Objects are cached by source mtime under xgnn_b200/lib/obj/ so rebuilds touch only what changed.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIBDIR = PKG / "lib"
OBJDIR = LIBDIR / "obj"
LIB = LIBDIR / "libxgnn_b200.so"

NVCC = os.environ.get("XG_NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
HOST_CXX = "/usr/bin/g++"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fopenmp,-Wall,-Wno-unknown-pragmas",
    "-ccbin", HOST_CXX,
    "--expt-relaxed-constexpr",
    "-DXG_BUILDING",
    f"-I{ROOT / 'include'}",
]


def _sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cc")))


def _headers_mtime() -> float:
    hs = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def _compile_one(src: Path, verbose: bool) -> Path:
    obj = OBJDIR / (src.name + ".o")
    newest = max(src.stat().st_mtime, _headers_mtime(), Path(__file__).stat().st_mtime)
    if obj.exists() and obj.stat().st_mtime >= newest:
        return obj
    cmd = [NVCC, *NVCC_FLAGS, "-x", "cu", "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"nvcc failed on {src.name}")
    if verbose:
        sys.stderr.write(r.stderr)
    return obj


def build_lib(verbose: bool = False, force: bool = False) -> Path:
    OBJDIR.mkdir(parents=True, exist_ok=True)
    if force:
        for o in OBJDIR.glob("*.o"):
            o.unlink()
    srcs = _sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose), srcs))
    if LIB.exists() and all(LIB.stat().st_mtime >= o.stat().st_mtime for o in objs):
        return LIB
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", HOST_CXX,
           "-Xcompiler", "-fPIC,-fopenmp", "-Xlinker", "--exclude-libs,ALL", "-Xlinker", f"--version-script={CSRC / 'exports.lds'}",
           "-o", str(LIB), *map(str, objs), "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    return LIB


def build_oracle() -> None:
    """oracle/liboracle.so always; oracle/_ref/* only where /root/reference exists."""
    subprocess.run(["make", "-s", "-C", str(ROOT / "oracle"), "oracle"], check=True)
    if Path("/root/reference/samgraph").is_dir():
        subprocess.run(["make", "-s", "-C", str(ROOT / "oracle"), "ref"], check=True)
        # the reference's CUDA samplers / ICS'22 solver / PreSampler compiled in place for sm_100 (second oracle, tests only)
        subprocess.run(["make", "-s", "-j8", "-C", str(ROOT / "oracle"), "refcuda"], check=True)


def build_all(verbose: bool = False, force: bool = False) -> Path:
    lib = build_lib(verbose=verbose, force=force)
    build_oracle()
    return lib


if __name__ == "__main__":
    p = build_all(verbose="-v" in sys.argv, force="-f" in sys.argv)
    print(p)
