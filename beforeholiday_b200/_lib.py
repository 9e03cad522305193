"""ctypes view of the C ABI (include/mt_b200.h) -- what a non-torch host (or the reference's maintainer writing a
cgo/JNI-style stub) would bind.  Used by the parity tests and bench.py to drive libmt_b200.so directly with raw
device pointers; torch is only the owner of the device memory here.

No compute happens in this file and there is no fallback: if the library is missing, loading raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_PKG = Path(__file__).resolve().parent
MT_MAX_DEPTH = 5
MT_F32, MT_F16, MT_BF16, MT_F64 = 0, 1, 2, 3

EXPORTS = [
    "mt_last_error", "mt_version", "mt_multi_tensor_scale", "mt_multi_tensor_axpby", "mt_norm_scratch_bytes",
    "mt_lamb_scratch_bytes", "mt_multi_tensor_l2norm", "mt_multi_tensor_l2norm_scale", "mt_multi_tensor_adam",
    "mt_multi_tensor_sgd", "mt_multi_tensor_lamb", "mt_multi_tensor_l2norm_mp", "mt_multi_tensor_lamb_mp",
    "mt_multi_tensor_adagrad", "mt_multi_tensor_novograd", "mt_multi_tensor_lars", "mt_multi_tensor_lamb_stage1", "mt_multi_tensor_lamb_stage2",
    "mt_multi_tensor_maybe_cast", "mt_dist_lamb_scratch_bytes", "mt_multi_tensor_lamb_compute_update_term", "mt_multi_tensor_lamb_update_weights",
    "mt_multi_tensor_fused_adam", "mt_multi_tensor_fused_adam_capturable", "mt_multi_tensor_fused_adam_p2p", "mt_multi_tensor_fused_adam_p2p_hybrid", "mt_multi_tensor_l2norm_noalloc", "mt_noalloc_resident_bytes",
    "mt_table_cache_stats", "mt_table_cache_clear", "mt_launch_count",
]


class MtTensorLists(C.Structure):
    _fields_ = [
        ("depth", C.c_int32),
        ("device", C.c_int32),
        ("ntensors", C.c_int64),
        ("ptrs", C.POINTER(C.c_void_p)),
        ("numel", C.POINTER(C.c_int64)),
        ("dtype", C.c_int32 * MT_MAX_DEPTH),
    ]


class MtCacheStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("hits", "misses", "uploads_bytes", "entries", "device_bytes",
                                           "evictions", "cross_stream_waits")]


_lib = None


def lib_path() -> Path:
    return _PKG / "libmt_b200.so"


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        p = lib_path()
        if not p.exists():
            raise OSError(f"{p} is missing: build it with `python -m beforeholiday_b200._build` (no CPU fallback exists)")
        L = C.CDLL(str(p))
        L.mt_last_error.restype = C.c_char_p
        L.mt_version.restype = C.c_char_p
        L.mt_norm_scratch_bytes.restype = C.c_size_t
        L.mt_lamb_scratch_bytes.restype = C.c_size_t
        L.mt_launch_count.restype = C.c_uint64
        L.mt_noalloc_resident_bytes.restype = C.c_size_t
        _lib = L
    return _lib


class MtError(RuntimeError):
    pass


def check(rc: int) -> None:
    if rc != 0:
        raise MtError(lib().mt_last_error().decode())


def _tag(dtype) -> int:
    import torch

    return {torch.float32: MT_F32, torch.float16: MT_F16, torch.bfloat16: MT_BF16, torch.float64: MT_F64}[dtype]


class Lists:
    """Owns the host arrays behind an MtTensorLists built from lists of torch CUDA tensors."""

    def __init__(self, tensor_lists):
        depth = len(tensor_lists)
        n = len(tensor_lists[0])
        self.ptr_arr = (C.c_void_p * (depth * n))()
        self.numel_arr = (C.c_int64 * n)()
        for d, l in enumerate(tensor_lists):
            assert len(l) == n
            for t, x in enumerate(l):
                self.ptr_arr[d * n + t] = x.data_ptr()
        for t, x in enumerate(tensor_lists[0]):
            self.numel_arr[t] = x.numel()
        self.tl = MtTensorLists()
        self.tl.depth = depth
        self.tl.device = tensor_lists[0][0].device.index or 0
        self.tl.ntensors = n
        self.tl.ptrs = C.cast(self.ptr_arr, C.POINTER(C.c_void_p))
        self.tl.numel = C.cast(self.numel_arr, C.POINTER(C.c_int64))
        for d, l in enumerate(tensor_lists):
            self.tl.dtype[d] = _tag(l[0].dtype)
        self.keep = tensor_lists

    @property
    def ref(self):
        return C.byref(self.tl)


def stream_ptr(device=None):
    import torch

    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def cache_stats(device: int = 0) -> dict:
    """Counters of the device tensor-table cache (hits, misses, uploads_bytes, entries, device_bytes)."""
    st = MtCacheStats()
    check(lib().mt_table_cache_stats(int(device), C.byref(st)))
    return {n: int(getattr(st, n)) for n, _ in MtCacheStats._fields_}
