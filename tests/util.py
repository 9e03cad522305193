"""Shared helpers for the parity tests."""
from __future__ import annotations

import torch

INT_VIEW = {torch.float32: torch.int32, torch.float16: torch.int16, torch.bfloat16: torch.int16,
            torch.float64: torch.int64}


def bits_equal(a: torch.Tensor, b: torch.Tensor) -> bool:
    """Bit-for-bit equality; NaNs compare equal to NaNs regardless of payload."""
    a = a.detach().cpu().contiguous()
    b = b.detach().cpu().contiguous()
    assert a.dtype == b.dtype and a.shape == b.shape, (a.dtype, b.dtype, a.shape, b.shape)
    same = a.view(INT_VIEW[a.dtype]) == b.view(INT_VIEW[b.dtype])
    both_nan = a.isnan() & b.isnan()
    return bool((same | both_nan).all())


def n_mismatch(a, b) -> int:
    a = a.detach().cpu().contiguous()
    b = b.detach().cpu().contiguous()
    same = (a.view(INT_VIEW[a.dtype]) == b.view(INT_VIEW[b.dtype])) | (a.isnan() & b.isnan())
    return int((~same).sum())


def ulp_diff_16(a: torch.Tensor, b: torch.Tensor) -> int:
    """Max distance in units of last place between two fp16/bf16 tensors (finite values)."""
    a = a.detach().cpu().contiguous()
    b = b.detach().cpu().contiguous()

    def key(x):
        i = x.view(torch.int16).to(torch.int32)
        return torch.where(i < 0, -(i & 0x7fff), i)  # sign-magnitude -> monotone integer

    return int((key(a) - key(b)).abs().max()) if a.numel() else 0


def rel_close(a, b, rtol, atol):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    return bool(((a - b).abs() <= rtol * b.abs() + atol).all())


def max_rel(a, b, floor=1e-30):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    return float(((a - b).abs() / b.abs().clamp_min(floor)).max())


def cpu_clone(lists):
    return [[t.detach().cpu().clone() for t in l] for l in lists]
