"""ctypes binding of libf5e_b200.so (the C ABI declared in include/f5e_b200.h).

There is deliberately no fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
_SUFFIX = os.environ.get("F5E_LIB_SUFFIX", "")  # developer knob, see build.py
_LIB_PATH = _HERE / (f"libf5e_b200_{_SUFFIX}.so" if _SUFFIX else "libf5e_b200.so")
_lib = None

ACT_NONE, ACT_GELU_TANH, ACT_GELU_ERF, ACT_MISH, ACT_RELU, ACT_SILU = range(6)


class Epilogue(C.Structure):
    """Mirror of `f5e_epilogue` (include/f5e_b200.h)."""

    _fields_ = [
        ("out", C.c_void_p), ("ldo", C.c_int32), ("out_f32", C.c_int32),
        ("bias", C.c_void_p), ("act", C.c_int32),
        ("gate", C.c_void_p), ("gate_stride", C.c_int32),
        ("residual", C.c_void_p), ("ldr", C.c_int32),
        ("row_len", C.c_void_p), ("n_pad", C.c_int32),
        ("rope_inner", C.c_int32), ("rope_heads", C.c_int32),
        ("residual_f32", C.c_int32), ("out_split3", C.c_int32), ("row_info", C.c_void_p),
        ("rope_inv_freq", C.c_void_p), ("ln_rstd", C.c_void_p), ("ln_a2", C.c_void_p),
        ("stats_out", C.c_void_p), ("stats_slots", C.c_int32),
    ]


class Guidance(C.Structure):
    """Mirror of `f5e_guidance` (include/f5e_b200.h)."""

    _fields_ = [("branches", C.c_int32), ("c0", C.c_float), ("c1", C.c_float)]


def lib_path() -> Path:
    return _LIB_PATH


def header_path() -> Path:
    return _HERE.parent / "include" / "f5e_b200.h"


def exported_symbols() -> list[str]:
    """Names of every function include/f5e_b200.h declares (the C ABI the library must export)."""
    import re

    src = re.sub(r"/\*.*?\*/", "", header_path().read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(f5e_[a-z0-9_]+)\s*\(", src)))


def load():
    """Load (building first if the sources are newer and nvcc is present) the shared library."""
    global _lib
    if _lib is not None:
        return _lib
    from . import build as _build

    build_error = None
    if os.environ.get("F5E_NO_BUILD") != "1":
        try:
            _build.build()
        except Exception as e:  # the prebuilt .so may still be usable (a box without nvcc); its stamp is checked below
            build_error = e
            if not _LIB_PATH.exists():
                raise RuntimeError(f"libf5e_b200.so is missing and could not be built: {e}") from e
    if not _LIB_PATH.exists():
        raise RuntimeError(f"{_LIB_PATH} not found; run `python -m f5e_tts_b200.build`")
    lib = C.CDLL(str(_LIB_PATH))
    lib.f5e_last_error.restype = C.c_char_p
    # A stale or half-built library must not run silently: its embedded stamp has to match the sources beside it
    # (developer variants built with F5E_LIB_SUFFIX / F5E_NVCC_FLAGS carry their own flags and are exempt).
    if not _SUFFIX and os.environ.get("F5E_ALLOW_STALE") != "1":
        try:
            lib.f5e_build_stamp.restype = C.c_char_p
            have = lib.f5e_build_stamp().decode()
        except AttributeError:
            have = "missing"
        want = _build.source_stamp()
        if have != want:
            raise RuntimeError(f"{_LIB_PATH} was built from other sources or flags (stamp {have[:12]} != {want[:12]})"
                               + (f" and rebuilding failed: {build_error}" if build_error else "")
                               + "; run `python -m f5e_tts_b200.build --force` (F5E_ALLOW_STALE=1 overrides)")
    lib.f5e_launch_count.restype = C.c_int64
    lib.f5e_prof_end.restype = C.c_int64
    lib.f5e_grn_static_scratch_floats.restype = C.c_int64
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().f5e_last_error().decode(errors="replace")
        raise RuntimeError(f"libf5e_b200 {what} failed ({rc}): {msg}")


def ptr(t) -> C.c_void_p:
    """Device pointer of a torch tensor (None -> NULL)."""
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr() -> C.c_void_p:
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_checked_devices: set = set()


def require_device() -> None:
    """The current CUDA device must be sm_100 (checked once per device: cudaGetDeviceProperties costs ~1 ms)."""
    import torch

    dev = torch.cuda.current_device() if torch.cuda.is_available() else -1
    if dev in _checked_devices:
        return
    check(load().f5e_check_device(), "f5e_check_device")
    _checked_devices.add(dev)
