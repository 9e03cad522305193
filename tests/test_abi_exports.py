"""CPU-side checks of the boundary: the C-ABI library loads without a GPU and exports every symbol that
include/mt_b200.h declares; error paths that need no device behave; the amp_C adaptor imports and exposes the
reference's operator names (/root/reference/csrc/amp_C_frontend.cpp:165-194)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "mt_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    syms = declared_symbols()
    assert len(syms) >= 24
    for s in syms:
        assert hasattr(L, s), f"libmt_b200.so does not export {s}"
    assert set(syms) == set(_lib.EXPORTS), set(syms) ^ set(_lib.EXPORTS)
    assert b"sm_100a" in L.mt_version()


def test_argument_errors_need_no_gpu():
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    tl = _lib.MtTensorLists()
    tl.depth = 3  # scale wants 2
    tl.ntensors = 1
    rc = L.mt_multi_tensor_scale(8192, None, C.byref(tl), C.c_float(1.0), None)
    assert rc == -1 and b"depth" in L.mt_last_error()
    assert L.mt_multi_tensor_scale(8192, None, None, C.c_float(1.0), None) == -1
    assert L.mt_norm_scratch_bytes(0, None) == 0
    assert L.mt_launch_count() == 0


def test_scratch_size_queries():
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    numel = (C.c_int64 * 3)(65536 * 2 + 1, 0, 5)
    ptrs = (C.c_void_p * 3)(256, 0, 512)
    tl = _lib.MtTensorLists()
    tl.depth, tl.ntensors, tl.device = 1, 3, 0
    tl.ptrs = C.cast(ptrs, C.POINTER(C.c_void_p))
    tl.numel = C.cast(numel, C.POINTER(C.c_int64))
    assert L.mt_norm_scratch_bytes(65536, C.byref(tl)) == 4 * (3 + 0 + 1)
    assert L.mt_norm_scratch_bytes(1 << 30, C.byref(tl)) == 4 * 2
    # element-wise families pick their own tile: 16384 elements once that yields >= 4 waves of 4 CTAs/SM (148 SMs assumed
    # without a device), shrinking to 8192 / 4096 / 2048 for smaller problems.  Sharded LAMB scratch = 2 floats per tile.
    L.mt_dist_lamb_scratch_bytes.restype = C.c_size_t
    for n0, tile in ((2048 * 3 + 1, 2048), (16384 * 3000, 16384), (8192 * 2400 + 5, 8192), (4096 * 2368, 4096)):
        numel[0], numel[1], numel[2] = n0, 0, 10
        tiles = -(-n0 // tile) + 0 + 1
        assert L.mt_dist_lamb_scratch_bytes(C.byref(tl)) == 2 * 4 * tiles, (n0, tile)


def test_amp_C_exports_reference_names():
    import beforeholiday_b200

    m = beforeholiday_b200.load_amp_C()
    for name in ("multi_tensor_scale", "multi_tensor_sgd", "multi_tensor_axpby", "multi_tensor_l2norm",
                 "multi_tensor_l2norm_scale", "multi_tensor_lamb_stage1_cuda", "multi_tensor_lamb_stage2_cuda",
                 "multi_tensor_adam", "multi_tensor_lamb", "multi_tensor_fused_adam", "multi_tensor_l2norm_mp",
                 "multi_tensor_lamb_mp", "multi_tensor_adagrad", "multi_tensor_novograd", "multi_tensor_lars",
                 "multi_tensor_fused_adam_p2p"):
        assert callable(getattr(m, name)), name


def test_cpu_tensor_is_rejected_like_the_reference():
    import torch

    import beforeholiday_b200

    m = beforeholiday_b200.load_amp_C()
    with pytest.raises(RuntimeError, match="expected input to be on cuda"):
        m.multi_tensor_scale(8192, torch.zeros(1, dtype=torch.int32), [[torch.zeros(4)], [torch.zeros(4)]], 1.0)


def test_install_registers_reference_module_names():
    import sys

    import beforeholiday_b200

    m = beforeholiday_b200.install()
    assert sys.modules["amp_C"] is m and sys.modules["distributed_adam_cuda"] is m
    from beforeholiday_b200.multi_tensor_apply import multi_tensor_applier, multi_tensor_applier_l2norm

    assert multi_tensor_applier.available and multi_tensor_applier.chunk_size == 256 * 32
    assert multi_tensor_applier_l2norm.chunk_size == 2048 * 32
    calls = []
    multi_tensor_applier(lambda *a: calls.append(a), "flag", [["t"]], 1.5, 2)
    assert calls == [(8192, "flag", [["t"]], 1.5, 2)]


def test_product_does_not_import_the_oracle():
    # the oracle is test infrastructure: nothing under beforeholiday_b200/ may import, include, link or load it
    # (comments may cite it)
    py = re.compile(r"^\s*(from|import)\s+oracle\b|libmt_oracle|CDLL\(.*oracle", re.M)
    cc = re.compile(r"#\s*include\s*[\"<][^\">]*oracle", re.M)
    for p in (ROOT / "beforeholiday_b200").rglob("*"):
        if p.suffix == ".py":
            assert not py.search(p.read_text()), p
        elif p.suffix in (".cu", ".cuh", ".cpp", ".h"):
            assert not cc.search(p.read_text()), p


def test_applier_forwards_with_its_chunk_size_and_reports_availability():
    from beforeholiday_b200.multi_tensor_apply import MultiTensorApply, multi_tensor_applier, multi_tensor_applier_l2norm

    assert MultiTensorApply.available and multi_tensor_applier.available and MultiTensorApply.import_err is None
    assert (multi_tensor_applier.chunk_size, multi_tensor_applier_l2norm.chunk_size) == (256 * 32, 2048 * 32)
    seen = []
    out = multi_tensor_applier(lambda *a: seen.append(a) or "ret", "flag", [["t"]], 1.5, True)
    assert out == "ret" and seen == [(8192, "flag", [["t"]], 1.5, True)]
    # unavailable extension: every call raises with the original import error in the message
    MultiTensorApply.available, MultiTensorApply.import_err = False, ImportError("no such file")
    try:
        with pytest.raises(RuntimeError, match="did not import.*no such file"):
            multi_tensor_applier(lambda *a: None, None, [[]])
    finally:
        MultiTensorApply.available, MultiTensorApply.import_err = True, None
