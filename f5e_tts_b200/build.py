"""Build libf5e_b200.so in-tree with nvcc for sm_100a (no torch extension machinery: the library is a
plain C-ABI shared object loaded with ctypes)."""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
# F5E_LIB_SUFFIX (developer knob): build / load libf5e_b200_<suffix>.so with F5E_NVCC_FLAGS beside the product library,
# so several compile-time variants of a kernel can travel to the GPU box in one snapshot (tools/attn_lab.py).
SUFFIX = os.environ.get("F5E_LIB_SUFFIX", "")
LIB = HERE / (f"libf5e_b200_{SUFFIX}.so" if SUFFIX else "libf5e_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
] + os.environ.get("F5E_NVCC_FLAGS", "").split()


def _sources():
    return sorted(CSRC.glob("*.cu"))


def _stamp() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
                    + [HERE.parent / "include" / "f5e_b200.h"]):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def source_stamp() -> str:
    """Hash of the sources + flags a current library was built from (compare with the library's f5e_build_stamp())."""
    return _stamp()


def build(force: bool = False, verbose: bool = False) -> Path:
    stamp_file = HERE / "csrc" / (".build_stamp" + ("_" + SUFFIX if SUFFIX else ""))
    stamp = _stamp()
    if not force and LIB.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return LIB
    if not Path(NVCC).exists():
        if LIB.exists():  # no compiler here: the shipped library is all there is; _lib.load() checks its embedded stamp
            return LIB
        raise RuntimeError(f"nvcc not found at {NVCC} and no prebuilt {LIB}")
    objdir = HERE / ("build_" + SUFFIX if SUFFIX else "build")
    objdir.mkdir(exist_ok=True)

    def compile_one(src: Path) -> Path:
        obj = objdir / (src.stem + ".o")
        cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)]
        if src.name == "c_api.cu":  # f5e_build_stamp(): lets the loader tell a stale library from a current one
            cmd.insert(1, f'-DF5E_BUILD_STAMP="{stamp}"')
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr, file=sys.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(compile_one, _sources()))
    cmd = [NVCC, "-shared", "-o", str(LIB), *map(str, objs), "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp_file.write_text(stamp)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
