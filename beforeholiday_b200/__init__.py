"""beforeholiday_b200 -- B200-native (sm_100a) multi-tensor-apply optimizer path.

Drop-in for the reference's hot path only:

    amp_C.multi_tensor_{scale,axpby,l2norm,l2norm_scale,adam,sgd,lamb,lamb_stage1_cuda,lamb_stage2_cuda}
    apex.multi_tensor_apply.multi_tensor_applier           -> beforeholiday_b200.multi_tensor_apply
    apex.optimizers.FusedAdam / FusedLAMB / FusedSGD        -> beforeholiday_b200.optimizers
    apex.amp.scaler.LossScaler                              -> beforeholiday_b200.amp.LossScaler
    apex.contrib.optimizers.DistributedFusedAdam (sharded)  -> beforeholiday_b200.contrib.optimizers

Everything computes in hand-written CUDA behind the C ABI of include/mt_b200.h (libmt_b200.so); there is
no CPU or PyTorch-eager fallback: importing `amp_C` fails loudly when the extension has not been built
(`python -m beforeholiday_b200._build`, or `__graft_entry__.build()`).
"""
from __future__ import annotations

import importlib
import sys

__version__ = "0.1.0"


def load_amp_C():
    """Import the compiled extension; raise (never fall back) if it is missing."""
    try:
        import torch  # noqa: F401  (libtorch must be loaded before the extension)

        return importlib.import_module("beforeholiday_b200.amp_C")
    except ImportError as e:  # pragma: no cover - exercised only on broken installs
        raise ImportError(
            "beforeholiday_b200.amp_C (the CUDA extension) is not built or failed to load: "
            f"{e}.  Build it in-tree with `python -m beforeholiday_b200._build`.  There is no CPU fallback."
        ) from e


def install(as_apex: bool = True):
    """Register the extension under the reference's module names so unmodified callers find it.

    After ``install()``: ``import amp_C``, ``import distributed_adam_cuda`` and ``import distributed_lamb_cuda`` resolve to this package's
    extension (the reference's optimizers do exactly those imports, e.g. apex/optimizers/fused_adam.py:75,
    apex/contrib/optimizers/distributed_fused_adam.py:291-293).
    """
    mod = load_amp_C()
    sys.modules.setdefault("amp_C", mod)
    if as_apex:
        sys.modules.setdefault("distributed_adam_cuda", mod)
        sys.modules.setdefault("distributed_lamb_cuda", mod)   # apex/contrib/optimizers/distributed_fused_lamb.py:99-106
        sys.modules.setdefault("fused_adam_cuda", mod)         # only maybe_cast_mt is carried over (:100, :856-861)
    return mod
