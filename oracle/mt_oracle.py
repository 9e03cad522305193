"""ctypes front-end of the CPU oracle (oracle/mt_oracle.c) on torch CPU tensors.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs (and tools/bench_ops.py, which extends those baseline legs to the other configs:
the CPU step timed beside a GPU number).  Nothing under beforeholiday_b200/ imports it.

The functions mirror the reference's ``amp_C`` operator signatures
(csrc/amp_C_frontend.cpp:165-194 in /root/reference) but run on CPU tensors, so a parity
test reads ``amp_C.multi_tensor_adam(...)`` on CUDA tensors next to
``oracle.multi_tensor_adam(...)`` on their CPU clones.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import torch

_HERE = Path(__file__).resolve().parent
_LIB = None

DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2, torch.float64: 3, torch.uint8: 4}   # uint8 = e5m2 packing


def build(force: bool = False) -> Path:
    so = _HERE / "libmt_oracle.so"
    src = _HERE / "mt_oracle.c"
    if force or not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "-s", "-B", "libmt_oracle.so"], check=True)
    return so


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(str(build()))
        _LIB.orc_scale.restype = C.c_int
        _LIB.orc_axpby.restype = C.c_int
        _LIB.orc_l2norm.restype = C.c_int
        _LIB.orc_num_threads.restype = C.c_int
    return _LIB


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _f(x):
    return C.c_float(float(x))


def _check_lists(tensor_lists, depth):
    assert len(tensor_lists) == depth, "tensor_lists.size() != depth"
    n0 = len(tensor_lists[0])
    assert n0 > 0
    for l in tensor_lists:
        assert len(l) == n0, "Size mismatch among tensor lists"
        for a, b in zip(l, tensor_lists[0]):
            assert a.device.type == "cpu" and a.is_contiguous() or a.is_contiguous(memory_format=torch.channels_last)
            assert a.numel() == b.numel(), "Size mismatch"


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(C.c_int(n))


def num_threads() -> int:
    return lib().orc_num_threads()


# ----------------------------------------------------------------------------------------------
def multi_tensor_scale(chunk_size, noop_flag, tensor_lists, scale):
    _check_lists(tensor_lists, 2)
    L = lib()
    for i, o in zip(*tensor_lists):
        f = L.orc_scale(_p(i), DT[i.dtype], _p(o), DT[o.dtype], C.c_int64(i.numel()), _f(scale))
        if f:
            noop_flag.fill_(1)


def multi_tensor_axpby(chunk_size, noop_flag, tensor_lists, a, b, arg_to_check):
    _check_lists(tensor_lists, 3)
    L = lib()
    for x, y, o in zip(*tensor_lists):
        f = L.orc_axpby(_p(x), DT[x.dtype], _p(y), DT[y.dtype], _p(o), DT[o.dtype], C.c_int64(x.numel()),
                        _f(a), _f(b), C.c_int(arg_to_check))
        if f:
            noop_flag.fill_(1)


def _l2norm(chunk_size, noop_flag, xs, per_tensor, outs=None, scale=0.0):
    L = lib()
    n = len(xs)
    ptrs = (C.c_void_p * n)(*[x.data_ptr() for x in xs])
    numel = (C.c_int64 * n)(*[x.numel() for x in xs])
    ret = torch.empty(1, dtype=torch.float32)
    rpt = torch.empty(n if per_tensor else 0, dtype=torch.float32)
    optrs = (C.c_void_p * n)(*[o.data_ptr() for o in outs]) if outs is not None else None
    f = L.orc_l2norm(ptrs, numel, C.c_int64(n), DT[xs[0].dtype], C.c_int(chunk_size),
                     _p(ret), _p(rpt) if per_tensor else C.c_void_p(0),
                     optrs, DT[outs[0].dtype] if outs is not None else 0, _f(scale))
    if f:
        noop_flag.fill_(1)
    return ret, rpt


def multi_tensor_l2norm(chunk_size, noop_flag, tensor_lists, per_tensor=False):
    _check_lists(tensor_lists, 1)
    return _l2norm(chunk_size, noop_flag, tensor_lists[0], bool(per_tensor))


def multi_tensor_l2norm_scale(chunk_size, noop_flag, tensor_lists, scale, per_tensor=False):
    _check_lists(tensor_lists, 2)
    return _l2norm(chunk_size, noop_flag, tensor_lists[0], bool(per_tensor), tensor_lists[1], scale)


def multi_tensor_adam(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, eps, step, mode, bias_correction,
                      weight_decay):
    _check_lists(tensor_lists, 4)
    L = lib()
    dt = DT[tensor_lists[0][0].dtype]
    for g, p, m, v in zip(*tensor_lists):
        L.orc_adam(_p(g), _p(p), _p(m), _p(v), dt, C.c_int64(g.numel()), _f(lr), _f(beta1), _f(beta2), _f(eps),
                   C.c_int(step), C.c_int(mode), C.c_int(bias_correction), _f(weight_decay))


def multi_tensor_sgd(chunk_size, noop_flag, tensor_lists, wd, momentum, dampening, lr, nesterov, first_run,
                     wd_after_momentum, scale):
    depth = len(tensor_lists)
    _check_lists(tensor_lists, depth)
    assert depth in (3, 4)
    L = lib()
    noop = (C.c_int * 1)(int(noop_flag.item()))
    for idx in range(len(tensor_lists[0])):
        g, w, mom = tensor_lists[0][idx], tensor_lists[1][idx], tensor_lists[2][idx]
        out = tensor_lists[3][idx] if depth == 4 else None
        if out is not None:
            assert out.dtype == torch.float16, "Additional output tensors should always be fp16."
        L.orc_sgd(_p(g), DT[g.dtype], _p(w), DT[w.dtype], _p(mom), _p(out), C.c_int64(g.numel()),
                  _f(wd), _f(momentum), _f(dampening), _f(lr), C.c_int(bool(nesterov)), C.c_int(bool(first_run)),
                  C.c_int(bool(wd_after_momentum)), _f(scale), noop)


def multi_tensor_lamb(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, eps, step, bias_correction,
                      weight_decay, grad_averaging, mode, global_grad_norm, max_grad_norm, use_nvlamb=False,
                      return_norms=False):
    _check_lists(tensor_lists, 4)
    L = lib()
    n = len(tensor_lists[0])
    arr = lambda l: (C.c_void_p * n)(*[t.data_ptr() for t in l])
    numel = (C.c_int64 * n)(*[t.numel() for t in tensor_lists[0]])
    pn = torch.empty(n, dtype=torch.float32)
    un = torch.empty(n, dtype=torch.float32)
    L.orc_lamb(arr(tensor_lists[0]), arr(tensor_lists[1]), arr(tensor_lists[2]), arr(tensor_lists[3]), numel,
               C.c_int64(n), DT[tensor_lists[0][0].dtype], C.c_int(chunk_size),
               _f(lr), _f(beta1), _f(beta2), _f(eps), C.c_int(step), C.c_int(bias_correction), _f(weight_decay),
               C.c_int(grad_averaging), C.c_int(mode), _f(float(global_grad_norm)), _f(max_grad_norm),
               C.c_int(bool(use_nvlamb)), _p(pn), _p(un))
    if return_norms:
        return pn, un


def multi_tensor_l2norm_mp(chunk_size, noop_flag, tensor_lists, per_tensor=False):
    """csrc/multi_tensor_l2norm_kernel_mp.cu: as multi_tensor_l2norm, but nothing is computed once the flag is set
    (the reference then returns uninitialised tensors; None here)."""
    if int(noop_flag.item()) != 0:
        return None, None
    ret, rpt = multi_tensor_l2norm(chunk_size, noop_flag, tensor_lists, per_tensor)
    if int(noop_flag.item()) != 0:   # cleanup also returns early when the main pass raised the flag (:121-123)
        return None, None
    return ret, rpt


def multi_tensor_lamb_mp(chunk_size, noop_flag, tensor_lists, lr, beta1, beta2, eps, step, bias_correction, weight_decay,
                         grad_averaging, mode, global_grad_norm, max_grad_norm, use_nvlamb, found_inf, inv_scale):
    """csrc/multi_tensor_lamb_mp.cu:367-496.  lr / step / norms / inv_scale are 1-element tensors."""
    depth = len(tensor_lists)
    assert depth in (4, 5)
    _check_lists(tensor_lists, depth)
    if int(noop_flag.item()) != 0:
        return
    L = lib()
    gs, ps, ms, vs = tensor_lists[:4]
    outs = tensor_lists[4] if depth == 5 else [None] * len(gs)
    g_dt, p_dt = DT[gs[0].dtype], DT[ps[0].dtype]
    beta3 = C.c_float(1.0 - C.c_float(beta1).value).value if grad_averaging == 1 else 1.0
    pn, _ = None, None
    _, pnorm = _l2norm(chunk_size, noop_flag, ps, True)
    if int(noop_flag.item()) != 0:
        return
    for g, p, m, v in zip(gs, ps, ms, vs):
        L.orc_lamb_mp_stage1(_p(g), g_dt, _p(p), _p(m), _p(v), p_dt, C.c_int64(g.numel()), _f(beta1), _f(beta2), _f(beta3),
                             C.c_int(int(step)), C.c_int(bias_correction), _f(eps), C.c_int(mode), _f(weight_decay),
                             _f(float(global_grad_norm)), _f(float(max_grad_norm)), _f(float(inv_scale)))
    _, unorm = _l2norm(chunk_size, noop_flag, gs, True)
    if int(noop_flag.item()) != 0:
        return
    for i, (g, p, o) in enumerate(zip(gs, ps, outs)):
        L.orc_lamb_mp_stage2(_p(g), g_dt, _p(p), p_dt, _p(o), C.c_int64(g.numel()), _f(float(lr)), _f(weight_decay),
                             C.c_int(bool(use_nvlamb)), _f(pnorm[i]), _f(unorm[i]))
    return pnorm, unorm


def multi_tensor_adagrad(chunk_size, noop_flag, tensor_lists, lr, epsilon, mode, weight_decay):
    _check_lists(tensor_lists, 3)
    L = lib()
    for g, p, h in zip(*tensor_lists):
        L.orc_adagrad(_p(g), _p(p), _p(h), DT[g.dtype], C.c_int64(g.numel()), _f(lr), _f(epsilon), C.c_int(mode),
                      _f(weight_decay))


def multi_tensor_novograd(chunk_size, noop_flag, tensor_lists, grad_norms, lr, beta1, beta2, epsilon, step,
                          bias_correction, weight_decay, grad_averaging, mode, norm_type):
    """grad_norms: float32 CPU tensor [ntensors], blended in place then used."""
    _check_lists(tensor_lists, 3)
    L = lib()
    L.orc_norm_blend.restype = C.c_int
    one_minus = C.c_float(1.0 - C.c_float(beta2).value).value
    for i, (g, p, m) in enumerate(zip(*tensor_lists)):
        out = (C.c_float * 1)(float(grad_norms[i]))
        if L.orc_norm_blend(_p(g), DT[g.dtype], C.c_int64(g.numel()), C.c_int(chunk_size), out, C.c_int(norm_type),
                            _f(beta2), _f(one_minus)):
            noop_flag.fill_(1)
        grad_norms[i] = out[0]
    for i, (g, p, m) in enumerate(zip(*tensor_lists)):
        L.orc_novograd(_p(g), _p(p), _p(m), DT[g.dtype], C.c_int64(g.numel()), _f(grad_norms[i]), _f(lr), _f(beta1),
                       _f(beta2), _f(epsilon), C.c_int(step), C.c_int(bias_correction), _f(weight_decay),
                       C.c_int(grad_averaging), C.c_int(mode))


def multi_tensor_lars(chunk_size, noop_flag, tensor_lists, grad_norms, param_norms, lr, trust_coefficient, epsilon,
                      weight_decay, momentum, dampening, nesterov, first_run, wd_after_momentum, scale, is_skipped):
    depth = len(tensor_lists)
    assert depth in (3, 4)
    _check_lists(tensor_lists, depth)
    L = lib()
    noop = (C.c_int * 1)(int(noop_flag.item()))
    for i in range(len(tensor_lists[0])):
        g, w, mom = tensor_lists[0][i], tensor_lists[1][i], tensor_lists[2][i]
        out = tensor_lists[3][i] if depth == 4 else None
        L.orc_lars(_p(g), DT[g.dtype], _p(w), DT[w.dtype], _p(mom), _p(out), C.c_int64(g.numel()), _f(grad_norms[i]),
                   _f(param_norms[i]), _f(lr), _f(trust_coefficient), _f(epsilon), _f(weight_decay), _f(momentum),
                   C.c_int(bool(nesterov)), C.c_int(bool(is_skipped)), noop)


def multi_tensor_lamb_stage1_cuda(chunk_size, noop_flag, tensor_lists, per_tensor_decay, step, beta1, beta2, eps,
                                  global_grad_norm, max_global_grad_norm):
    _check_lists(tensor_lists, 5)
    L = lib()
    for i, (g, p, m, v, u) in enumerate(zip(*tensor_lists)):
        L.orc_lamb_stage1_legacy(_p(g), DT[g.dtype], _p(p), _p(m), _p(v), DT[p.dtype], _p(u), DT[u.dtype],
                                 C.c_int64(g.numel()), _f(per_tensor_decay[i]), C.c_int(step), _f(beta1), _f(beta2),
                                 _f(eps), _f(float(global_grad_norm)), _f(max_global_grad_norm))


def multi_tensor_lamb_stage2_cuda(chunk_size, noop_flag, tensor_lists, per_tensor_param_norm,
                                  per_tensor_update_norm, lr, weight_decay, use_nvlamb=False):
    _check_lists(tensor_lists, 2)
    L = lib()
    for i, (p, u) in enumerate(zip(*tensor_lists)):
        L.orc_lamb_stage2_legacy(_p(p), DT[p.dtype], _p(u), DT[u.dtype], C.c_int64(p.numel()), _f(lr),
                                 _f(weight_decay), C.c_int(bool(use_nvlamb)), _f(per_tensor_param_norm[i]),
                                 _f(per_tensor_update_norm[i]))


def multi_tensor_fused_adam(chunk_size, noop_flag, tensor_lists, per_tensor_beta1, per_tensor_beta2,
                            per_tensor_bias_correction, per_tensor_eps, per_tensor_weight_decay, lr, grad_scale,
                            step, mode, honour_noop=False):
    """distributed_adam_cuda.multi_tensor_fused_adam (lists: p, m, v, g[, p_copy])."""
    depth = len(tensor_lists)
    assert depth in (4, 5)
    _check_lists(tensor_lists, depth)
    L = lib()
    noop = (C.c_int * 1)(int(noop_flag.item())) if honour_noop else None
    for i in range(len(tensor_lists[0])):
        p, m, v, g = (tensor_lists[k][i] for k in range(4))
        pc = tensor_lists[4][i] if depth == 5 else None
        assert p.dtype == m.dtype == v.dtype == torch.float32
        L.orc_dist_adam(_p(p), _p(m), _p(v), _p(g), DT[g.dtype], _p(pc), C.c_int64(p.numel()),
                        _f(per_tensor_beta1[i]), _f(per_tensor_beta2[i]), C.c_int(int(per_tensor_bias_correction[i])),
                        _f(per_tensor_eps[i]), _f(per_tensor_weight_decay[i]), _f(lr), _f(grad_scale),
                        C.c_int(step), C.c_int(mode), noop)


def convert(src: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    out = torch.empty(src.shape, dtype=dtype)
    lib().orc_convert(_p(src), DT[src.dtype], _p(out), DT[dtype], C.c_int64(src.numel()))
    return out


# ---- sharded LAMB (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:3-29) ----------------------------------
def multi_tensor_lamb_compute_update_term(chunk_size, noop_flag, tensor_lists, per_tensor_beta1, per_tensor_beta2,
                                          per_tensor_beta3, per_tensor_bias_correction, step, per_tensor_epsilon, mode,
                                          per_tensor_decay, global_scale, global_grad_norm, max_grad_norm,
                                          model_dtype=None):
    """Lists [g, p, m, v, u] of CPU tensors (p, m, v, u float32).  Returns a float64 [2, ntensors] tensor of the
    per-fragment sums of round_model(p)^2 and u^2 (the product's norm extension is checked against it)."""
    _check_lists(tensor_lists, 5)
    n = len(tensor_lists[0])
    sq = torch.zeros(2, n, dtype=torch.float64)
    if int(noop_flag.item()) != 0:
        return sq
    L = lib()
    mdt = DT[model_dtype] if model_dtype is not None else 0
    for i, (g, p, m, v, u) in enumerate(zip(*tensor_lists)):
        assert p.dtype == m.dtype == v.dtype == u.dtype == torch.float32
        out = (C.c_double * 2)(0.0, 0.0)
        L.orc_dist_lamb_stage1(_p(g), DT[g.dtype], _p(p), _p(m), _p(v), _p(u), C.c_int64(g.numel()),
                               _f(per_tensor_beta1[i]), _f(per_tensor_beta2[i]), C.c_int(int(per_tensor_bias_correction[i])),
                               C.c_int(int(step)), _f(per_tensor_epsilon[i]), C.c_int(mode), _f(per_tensor_decay[i]),
                               _f(float(global_scale)), _f(float(global_grad_norm)), _f(max_grad_norm), C.c_int(mdt), out)
        sq[0, i], sq[1, i] = out[0], out[1]
    return sq


def multi_tensor_lamb_update_weights(chunk_size, noop_flag, tensor_lists, per_tensor_param_norm, per_tensor_update_norm,
                                     update_norm_offset, learning_rate, per_tensor_decay, global_grad_norm, use_nvlamb):
    """Lists [u, p, p_copy] of CPU tensors (u, p float32)."""
    _check_lists(tensor_lists, 3)
    if int(noop_flag.item()) != 0:
        return
    L = lib()
    for i, (u, p, c) in enumerate(zip(*tensor_lists)):
        L.orc_dist_lamb_stage2(_p(u), _p(p), _p(c), DT[c.dtype], C.c_int64(u.numel()), _f(float(learning_rate)),
                               _f(per_tensor_decay[i]), C.c_int(bool(use_nvlamb)), _f(per_tensor_param_norm[i]),
                               _f(per_tensor_update_norm[int(update_norm_offset[i])]))


def maybe_cast_mt(chunk_size, overflow_flag, tensor_lists):
    """apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:734-783: converting copy, skipped when the flag is set."""
    _check_lists(tensor_lists, 2)
    if int(overflow_flag.item()) != 0:
        return
    for a, b in zip(*tensor_lists):
        b.copy_(a.to(b.dtype))
