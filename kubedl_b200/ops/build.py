"""In-tree build of libkubedl_b200.so (nvcc, sm_100a only).

The library is plain C ABI (include/kubedl_b200.h) so it needs neither torch headers nor
pybind11: each .cu compiles in seconds and the resulting .so travels with the repo snapshot
to the GPU box.  A content hash of the sources is stored next to the .so; `ensure_built()`
rebuilds only when the hash changed or the .so is missing.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIB_DIR = ROOT / "kubedl_b200" / "_lib"
LIB_PATH = LIB_DIR / "libkubedl_b200.so"
HASH_PATH = LIB_DIR / "libkubedl_b200.hash"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=default", "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build libkubedl_b200.so")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _source_hash() -> str:
    h = hashlib.sha256()
    files = sources() + sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h"))
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(ARCH_FLAGS + NVCC_FLAGS).encode())
    return h.hexdigest()


def is_current() -> bool:
    return LIB_PATH.exists() and HASH_PATH.exists() and HASH_PATH.read_text().strip() == _source_hash()


def build(verbose: bool = False, force: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link them into one shared library."""
    if not force and is_current():
        return LIB_PATH
    nvcc = _nvcc()
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    obj_dir = LIB_DIR / "obj"
    obj_dir.mkdir(exist_ok=True)

    def compile_one(src: Path) -> Path:
        obj = obj_dir / (src.stem + ".o")
        cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC)]
        if verbose:
            cmd += ["-Xptxas", "-v"]
        cmd += ["-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode != 0:
            sys.stderr.write(f"$ {' '.join(cmd)}\n{r.stdout}{r.stderr}\n")
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    tmp = LIB_PATH.with_suffix(".so.tmp")
    cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", str(tmp), *map(str, objs), "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stderr}")
    os.replace(tmp, LIB_PATH)
    HASH_PATH.write_text(_source_hash())
    return LIB_PATH


def ensure_built() -> Path:
    if is_current():
        return LIB_PATH
    if LIB_PATH.exists() and os.environ.get("KUBEDL_B200_NO_REBUILD") == "1":
        return LIB_PATH
    return build()


if __name__ == "__main__":
    p = build(verbose="-v" in sys.argv, force=True)
    print(p)
