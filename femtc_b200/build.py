"""Build libfemtc_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python femtc_b200/build.py [--force] [--verbose]      (run by path: importing the package needs the library)

Objects are cached under femtc_b200/csrc/_obj and rebuilt when a source or header is newer.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
# development knobs: FEMTC_LIB_SUFFIX=_x FEMTC_EXTRA_FLAGS="-DFOO=1" builds libfemtc_b200_x.so next to the
# product library (own object cache) for A/B timing of compile-time variants (tools/quick_bench.py --lib)
SUFFIX = os.environ.get("FEMTC_LIB_SUFFIX", "")
OBJ = CSRC / ("_obj" + SUFFIX)
LIB = PKG / f"libfemtc_b200{SUFFIX}.so"
SOURCES = ["api.cu", "generic_kernels.cu", "tiled_kernels.cu", "column_kernels.cu", "cell_kernels.cu", "tc_kernels.cu", "tf32_kernels.cu", "tmm_kernels.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-std=c++17", "-O3", "-lineinfo",
    "-Xcompiler", "-fPIC,-Wall,-Wextra",
    "--expt-relaxed-constexpr",
] + os.environ.get("FEMTC_EXTRA_FLAGS", "").split()


def _headers():
    return list(CSRC.glob("*.cuh")) + list((PKG.parent / "include").glob("*.h"))


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(exist_ok=True)
    hdrs = _headers()

    def compile_one(src: str):
        s = CSRC / src
        o = OBJ / (s.stem + ".o")
        if not force and not _stale(o, [s] + hdrs):
            return o, ""
        cmd = [NVCC, *FLAGS, "-Xptxas", "-v", "-c", str(s), "-o", str(o)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        (OBJ / (s.stem + ".ptxas.log")).write_text(r.stderr)
        return o, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        results = list(ex.map(compile_one, SOURCES))
    objs = [str(o) for o, _ in results]
    if force or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", str(LIB), *objs, "-gencode", "arch=compute_100a,code=sm_100a",
               "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        for _, log in results:
            if log:
                print(log)
    return LIB


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
