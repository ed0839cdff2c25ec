"""ctypes binding of libmingpt_b200.so (the C ABI declared in include/mingpt_b200.h).

The shared library is built in-tree by csrc/build.sh (nvcc, sm_100a). There is no fallback: if the
library is missing and cannot be built, importing any op raises.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import shutil
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmingpt_b200.so")
BUILD_SCRIPT = os.path.join(_HERE, "csrc", "build.sh")

_lib = None
_lock = threading.Lock()


class GemmArgs(C.Structure):
    """mirror of mgb_gemm_args (include/mingpt_b200.h)"""

    _fields_ = [
        ("a", C.c_void_p), ("lda", C.c_int64), ("a_mn_major", C.c_int32),
        ("b", C.c_void_p), ("ldb", C.c_int64), ("b_mn_major", C.c_int32),
        ("d", C.c_void_p), ("ldd", C.c_int64), ("d_f32", C.c_int32), ("accumulate", C.c_int32),
        ("m", C.c_int64), ("n", C.c_int64), ("k", C.c_int64),
        ("bias", C.c_void_p),
        ("residual", C.c_void_p), ("ldr", C.c_int64),
        ("act", C.c_int32),
        ("preact", C.c_void_p), ("ldp", C.c_int64),
        ("gate", C.c_void_p), ("ldg", C.c_int64), ("gate_kind", C.c_int32),
        ("split_k", C.c_int32),
        ("block_n", C.c_int32),
        ("alpha", C.c_float),
        ("colsum", C.c_void_p),
        ("rowdot_x", C.c_void_p), ("ld_rowdot", C.c_int64),
        ("rowdot_out", C.c_void_p), ("rowdot_t", C.c_int64),
        ("drop_seed", C.c_void_p), ("drop_p", C.c_float), ("drop_site", C.c_uint32),
    ]


class AdamWChunk(C.Structure):
    """mirror of mgb_adamw_chunk"""

    _fields_ = [("offset", C.c_int64), ("numel", C.c_int32), ("weight_decay", C.c_float)]


_P, _I64, _I32, _F = C.c_void_p, C.c_int64, C.c_int32, C.c_float

# name -> argtypes; every function returns int except the three introspection calls
SIGNATURES = {
    "mgb_gemm_bf16": [C.POINTER(GemmArgs), _P],
    "mgb_layernorm_fwd": [_P, _P, _P, _P, _P, _P, _I64, _I32, _F, _P],
    "mgb_layernorm_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I32, _P],
    "mgb_layernorm_bwd_colsum": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _I32, _P],
    "mgb_embedding_fwd": [_P, _P, _P, _P, _I64, _I64, _I32, _I64, _P, _P],
    "mgb_embedding_bwd": [_P, _P, _P, _P, _I64, _I64, _I32, _I64, _P],
    "mgb_token_windows": [_P, _I64, _P, _I64, _I32, _P, _P, _P, _P],
    "mgb_attention_fwd": [_P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _P],
    "mgb_attention_bwd": [_P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _P],
    "mgb_attention_bwd_colsum": [_P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _P],
    "mgb_attention_bwd_noalloc": [_P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _P],
    "mgb_attention_decode": [_P, _P, _I32, _I32, _I32, _I32, _I32, _F, _P],
    "mgb_dropout_bf16": [_P, _P, _P, _I64, _F, _P, C.c_uint32, _P],
    "mgb_dropout_bf16_colsum": [_P, _P, _I64, _I32, _F, _P, C.c_uint32, _P, _P],
    "mgb_attention_fwd_dropout": [_P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _F, _P, C.c_uint32, _P],
    "mgb_attention_bwd_dropout": [_P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _F, _P, C.c_uint32, _P],
    "mgb_attention_dropout_mask": [_P, _I32, _I32, _I32, _F, _P, C.c_uint32, _P],
    "mgb_attention_dropout_bits_words": [_I32, _I32, _I32, _P],
    "mgb_attention_dropout_bits": [_P, _I32, _I32, _I32, _I32, _F, _P, C.c_uint32, _P],
    "mgb_attention_fwd_dropout_bits": [_P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _F, _P, _P],
    "mgb_attention_bwd_dropout_bits": [_P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _I32, _I32, _F, _F, _P, _P, _P],
    "mgb_cross_entropy_fwd_bwd": [_P, _I64, _P, _P, _P, _I64, _I64, _F, _I32, _P, _P],
    "mgb_count_valid_targets": [_P, _I64, _P, _P],
    "mgb_loss_finalize": [_P, _I64, _P, _P, _P],
    "mgb_colsum_bf16": [_P, _I64, _I64, _I64, _P, _P],
    "mgb_act_bwd_bf16": [_P, _P, _P, _I64, _I32, _P],
    "mgb_act_fwd_bf16": [_P, _P, _P, _I64, _I32, _P],
    "mgb_cast_f32_to_bf16": [_P, _P, _I64, _P],
    "mgb_cast_bf16_to_f32": [_P, _P, _I64, _P],
    "mgb_axpy_f32": [_P, _P, _F, _I64, _P],
    "mgb_axpy_f32_dev": [_P, _P, _P, _I64, _P],
    "mgb_scale_bf16_dev": [_P, _I64, _P, _P],
    "mgb_sumsq_f32": [_P, _I64, _P, _P],
    "mgb_adamw_multi": [_P, _P, _P, _P, _P, _P, _I32, _F, _F, _F, _F, _I32, _F, _P, _F, _P, _P],
    "mgb_clip_scale_f32": [_P, _I64, _P, _F, _P],
    "mgb_set_option": [C.c_char_p, _I32],
}
INTROSPECTION = {"mgb_last_error": C.c_char_p, "mgb_version": C.c_int, "mgb_launch_count": C.c_int64}


STAMP_PATH = LIB_PATH + ".srchash"


def source_hash() -> str:
    """sha256 over the CUDA sources, the public header and the build script: what LIB_PATH must be built from"""
    src_dir = os.path.join(_HERE, "csrc")
    inc = os.path.join(os.path.dirname(_HERE), "include", "mingpt_b200.h")
    files = sorted(os.path.join(src_dir, f) for f in os.listdir(src_dir) if f.endswith((".cu", ".cuh", ".sh"))) + [inc]
    h = hashlib.sha256()       # == `cat <files> | sha256sum`, which is how csrc/build.sh writes the stamp
    for f in files:
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def build(verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a into LIB_PATH (nvcc cross-compiles without a GPU) and record the
    hash of the sources it was built from next to it."""
    import fcntl
    with open(LIB_PATH + ".lock", "w") as lock:       # several ranks may reach this at once
        fcntl.flock(lock, fcntl.LOCK_EX)
        if os.path.exists(LIB_PATH) and not _lib_is_stale():
            return LIB_PATH                           # another process built it while we waited
        res = subprocess.run(["bash", BUILD_SCRIPT], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"building {LIB_PATH} failed:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stdout.strip())
    return LIB_PATH


def _lib_is_stale() -> bool:
    """True when LIB_PATH was built from different sources than the ones on disk (content hash, not mtimes:
    a snapshot copy does not preserve modification times)."""
    if not os.path.exists(LIB_PATH):
        return True
    try:
        with open(STAMP_PATH) as f:
            return f.read().strip() != source_hash()
    except OSError:
        return True


def lib() -> C.CDLL:
    """Load (building first if the .so is absent) and type the library. Raises if unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        override = os.environ.get("MGB_LIB_PATH")      # A/B measurements: load another build of the library as is
        if override:
            handle = C.CDLL(override)
            for table, default in ((SIGNATURES, C.c_int), (INTROSPECTION, None)):
                for name, spec in table.items():
                    fn = getattr(handle, name, None)
                    if fn is None:
                        continue                             # an older build lacks the newer entry points
                    fn.argtypes = spec if default is not None else []
                    fn.restype = default if default is not None else spec
            _lib = handle
            return _lib
        if _lib_is_stale():
            # an edited .cu / .cuh / header must never silently run the old binary: rebuild, or refuse to load
            if shutil.which(os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")) is None:
                raise RuntimeError(f"{LIB_PATH} is missing or was built from other sources than the ones on disk, "
                                   "and nvcc is not available to rebuild it")
            build()
        handle = C.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.argtypes = argtypes
            fn.restype = C.c_int
        for name, restype in INTROSPECTION.items():
            fn = getattr(handle, name)
            fn.argtypes = []
            fn.restype = restype
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().mgb_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"mingpt_b200 {what} failed (code {rc}): {msg}")


def set_option(key: str, value: int) -> int:
    """process-wide tuning knob (see mgb_set_option); returns the previous value"""
    prev = lib().mgb_set_option(key.encode(), int(value))
    if prev < 0:
        raise KeyError(key)
    return prev


def launch_count() -> int:
    return int(lib().mgb_launch_count())
