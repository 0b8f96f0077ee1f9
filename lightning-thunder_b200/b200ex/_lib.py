"""ctypes binding of the C-ABI kernel library ``libb200k.so`` (declared in ``include/b200k.h``).

The library is the only compute backend of this package.  If it cannot be loaded the import of
any b200ex operation fails loudly with :class:`B200LibraryError`; there is no CPU or eager fallback.
"""

from __future__ import annotations

import ctypes
import os
import threading

__all__ = ["lib", "check", "B200LibraryError", "B200KernelError", "library_path", "EXPORTS"]

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(os.path.dirname(_HERE), "csrc", "libb200k.so")

c_void_p, c_int, c_int64, c_float, c_char_p = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_char_p


class B200LibraryError(ImportError):
    """libb200k.so is missing or does not export the expected ABI."""


class B200KernelError(RuntimeError):
    """A b200k entry point returned a non-zero status."""


# name -> (restype, argtypes); must list every symbol include/b200k.h declares.
EXPORTS: dict[str, tuple] = {
    "b200k_version": (c_int, []),
    "b200k_error_string": (c_char_p, [c_int]),
    "b200k_sm_count": (c_int, []),
    "b200k_rmsnorm_fwd": (c_int, [c_void_p] * 6 + [c_int64, c_int64, c_float, c_int, c_void_p]),
    "b200k_rmsnorm_bwd_scratch_bytes": (c_int64, [c_int64, c_int64]),
    "b200k_rmsnorm_bwd": (c_int, [c_void_p] * 8 + [c_int64, c_int64, c_int64, c_int, c_void_p]),
    "b200k_rmsnorm_bwd_noalloc": (c_int, [c_void_p] * 7 + [c_int64, c_int64, c_int, c_void_p]),
    "b200k_swiglu_fwd": (c_int, [c_void_p] * 3 + [c_int64, c_int, c_void_p]),
    "b200k_swiglu_bwd": (c_int, [c_void_p] * 5 + [c_int64, c_int, c_void_p]),
    "b200k_cross_entropy_fwd": (c_int, [c_void_p] * 7 + [c_int64, c_int64, c_int64, c_int, c_int, c_void_p]),
    "b200k_cross_entropy_bwd": (c_int, [c_void_p] * 7 + [c_int64, c_int64, c_int64, c_int, c_int, c_void_p]),
    "b200k_qkv_rope_fwd": (c_int, [c_void_p] * 6 + [c_int64, c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200k_qkv_rope_bwd": (c_int, [c_void_p] * 6 + [c_int64, c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200k_bias_act_fwd": (c_int, [c_void_p] * 4 + [c_int64, c_int64, c_int, c_int, c_void_p]),
    "b200k_bias_act_bwd_scratch_bytes": (c_int64, [c_int64, c_int64]),
    "b200k_bias_act_bwd": (c_int, [c_void_p] * 6 + [c_int64, c_int64, c_int64, c_int, c_int, c_void_p]),
    "b200k_layernorm_fwd": (c_int, [c_void_p] * 6 + [c_int64, c_int64, c_float, c_int, c_void_p]),
    "b200k_layernorm_bwd_scratch_bytes": (c_int64, [c_int64, c_int64]),
    "b200k_layernorm_bwd": (c_int, [c_void_p] * 9 + [c_int64, c_int64, c_int64, c_int, c_void_p]),
    "b200k_adamw_chunk_elems": (c_int64, [c_int]),
    "b200k_adamw": (c_int, [c_void_p, c_void_p, c_void_p, c_int64] + [ctypes.c_double] * 5 + [c_float, c_float, c_int, c_int, c_void_p]),
    "b200k_softmax_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "b200k_softmax_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "b200k_uniform_philox": (c_int, [c_void_p, c_int64, ctypes.c_uint64, ctypes.c_uint64, c_float, c_float, c_int, c_void_p]),
    "b200k_colsum": (c_int, [c_void_p] * 3 + [c_int64, c_int64, c_int64, c_int, c_void_p]),
    "b200k_pointwise": (c_int, [c_void_p, c_void_p]),
}

ABI_VERSION = 6

_lock = threading.Lock()
_lib = None


def library_path() -> str:
    return os.environ.get("B200K_LIBRARY", _LIB_PATH)


def lib() -> ctypes.CDLL:
    """Load (once, thread-safe) and return the kernel library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = library_path()
        if not os.path.exists(path):
            raise B200LibraryError(
                f"b200ex: kernel library not found at {path}. Build it with `make -C lightning-thunder_b200/csrc` "
                "(or __graft_entry__.build()). b200ex has no fallback path."
            )
        try:
            handle = ctypes.CDLL(path)
        except OSError as e:  # pragma: no cover - depends on the host
            raise B200LibraryError(f"b200ex: cannot load {path}: {e}") from e
        for name, (restype, argtypes) in EXPORTS.items():
            try:
                fn = getattr(handle, name)
            except AttributeError as e:
                raise B200LibraryError(f"b200ex: {path} does not export {name}; rebuild the library") from e
            fn.restype = restype
            fn.argtypes = argtypes
        if handle.b200k_version() != ABI_VERSION:
            raise B200LibraryError(f"b200ex: {path} has ABI {handle.b200k_version()}, expected {ABI_VERSION}; rebuild")
        _lib = handle
    return _lib


def check(status: int, what: str) -> None:
    """Raise :class:`B200KernelError` for a non-zero C-ABI status."""
    if status != 0:
        msg = lib().b200k_error_string(status).decode()
        raise B200KernelError(f"{what} failed with status {status}: {msg}")
