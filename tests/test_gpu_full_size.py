"""Parity at BASELINE.json's FULL sizes (configs C2-C5 of SURVEY.md section 8d), through properties that do not need a
second full copy of the state:

  * sampled bit-exact parity -- slices of the inputs (tensor heads, tails, tile boundaries, whole small tensors) are
    copied to the host BEFORE the op, pushed through the CPU oracle, and compared bit for bit with the same slices of
    the device result AFTER the op (the update is element-wise, so a slice's result does not depend on its position);
  * checksum of checksums -- the global norm equals the root of the summed squared per-tensor norms and matches a
    double-precision torch norm; per-tensor norms of sampled tensors match torch;
  * exact round trips -- scaling by 2^k and by 2^-k returns the input bits;
  * idempotence under the flag -- with the overflow flag raised the step leaves an integer checksum of p/m/v unchanged.

C2 (GPT-2-small, 124 M) and C4 (BERT-large, 336 M) are small enough to also run the oracle over EVERYTHING.
"""
import pytest
import torch

import workloads
from util import bits_equal, n_mismatch

pytestmark = pytest.mark.gpu

F32, F16, BF16 = torch.float32, torch.float16, torch.bfloat16


def _need_gb(gb):
    free, _ = torch.cuda.mem_get_info()
    if free < gb * 2 ** 30:
        pytest.skip(f"needs {gb} GB of free device memory, {free / 2 ** 30:.0f} GB available")


def _carve(flat, numels):
    out, o = [], 0
    for n in numels:
        out.append(flat[o:o + n])
        o += n
    return out


def _bit_checksum(t):
    """Order-independent integer checksum of the bit patterns (int64 sum of the 32-bit words)."""
    w = t.view(torch.int32) if t.element_size() == 4 else t.view(torch.int16)
    total = 0
    for i in range(0, w.numel(), 1 << 28):
        total += int(w[i:i + (1 << 28)].sum(dtype=torch.int64))
    return total


def _sample_slices(numels, tile=16384, width=4096, extra=()):
    """(tensor index, lo, hi) triples: head, first tile boundary, an interior tile boundary, tail of selected tensors."""
    order = sorted(range(len(numels)), key=lambda i: numels[i])
    picks = {0, len(numels) - 1, order[0], order[-1], order[len(order) // 2], *extra}
    out = []
    for t in sorted(picks):
        n = numels[t]
        spans = [(0, min(n, width)), (max(0, n - width), n)]
        for b in (tile, (n // tile // 2) * tile, (n // tile) * tile):
            if 8 <= b < n:
                spans.append((max(0, b - 9), min(n, b + 9)))
        for lo, hi in spans:
            if hi > lo:
                out.append((t, lo, hi))
    return out


# ------------------------------------------------------------------------------------------------ C5: Llama-3-8B Adam
@pytest.fixture(scope="module")
def llama_state():
    _need_gb(140)
    numels = workloads.numels(workloads.llama3_8b())
    P = sum(numels)
    assert P == 8030261248 and len(numels) == 291
    gen = torch.Generator(device="cuda").manual_seed(9876)
    p = torch.empty(P, dtype=F32, device="cuda").normal_(0, 0.02, generator=gen)
    m = torch.empty(P, dtype=F32, device="cuda").normal_(0, 1e-3, generator=gen)
    v = torch.empty(P, dtype=F32, device="cuda").normal_(0, 1e-3, generator=gen).square_()
    g = torch.empty(P, dtype=BF16, device="cuda").normal_(0, 1.0, generator=gen)
    pc = torch.zeros(P, dtype=BF16, device="cuda")
    yield dict(numels=numels, P=P, flat=dict(p=p, m=m, v=v, g=g, pc=pc),
               lists=[_carve(x, numels) for x in (p, m, v, g, pc)])
    del p, m, v, g, pc
    torch.cuda.empty_cache()


def test_llama3_8b_adam_step_sampled_parity_and_flag_idempotence(amp_C, oracle, llama_state):
    S = llama_state
    numels, lists = S["numels"], S["lists"]
    nt = len(numels)
    T = lambda val, dt=F32: torch.full((nt,), val, dtype=dt, device="cuda")
    wd = torch.tensor([0.1 if n > 4096 else 0.0 for n in numels], device="cuda")     # norms carry no decay
    hp = (T(0.9), T(0.95), T(1, torch.int32), T(1e-8), wd)
    # 1. flag raised: the whole 8 B-parameter step must be a no-op
    before = [_bit_checksum(S["flat"][k]) for k in ("p", "m", "v", "pc")]
    one = torch.ones(1, dtype=torch.int32, device="cuda")
    amp_C.multi_tensor_fused_adam(8192, one, lists, *hp, 1e-4, 1024.0, 7, 1, True)
    assert [_bit_checksum(S["flat"][k]) for k in ("p", "m", "v", "pc")] == before
    # 2. sampled bit-exact parity of a real step
    slices = _sample_slices(numels, extra=(1, 2, 145))
    host_in = [[lists[d][t][lo:hi].cpu().clone() for (t, lo, hi) in slices] for d in range(5)]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    launches = amp_C.launch_count()
    amp_C.multi_tensor_fused_adam(8192, flag, lists, *hp, 1e-4, 1024.0, 7, 1, True)
    assert amp_C.launch_count() - launches == 1          # 291 tensors, 8 G elements: ONE launch
    wd_h = wd.cpu()
    k = len(slices)
    oracle.multi_tensor_fused_adam(8192, torch.zeros(1, dtype=torch.int32), host_in, [0.9] * k, [0.95] * k, [1] * k,
                                   [1e-8] * k, [float(wd_h[t]) for (t, _, _) in slices], 1e-4, 1024.0, 7, 1)
    for d in (0, 1, 2, 4):
        for (t, lo, hi), want in zip(slices, host_in[d]):
            got = lists[d][t][lo:hi]
            assert bits_equal(got, want), f"list {d} tensor {t} [{lo},{hi}): {n_mismatch(got, want)} mismatches"
    # 3. the bf16 model copy is the rounded master (whole tensors: the 525 M-element embedding, a mid tensor, the head)
    for t in (0, 17, nt - 1):
        assert torch.equal(lists[4][t], lists[0][t].to(BF16))


def test_llama3_8b_norm_checksums_and_scale_round_trip(amp_C, llama_state):
    S = llama_state
    numels, g = S["numels"], S["lists"][3]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    tot, per = amp_C.multi_tensor_l2norm(65536, flag, [g], True)
    assert flag.item() == 0
    # checksum of checksums: global^2 == sum of per-tensor^2, and both agree with double-precision torch sums
    assert abs(float(tot) ** 2 - float((per.double() ** 2).sum())) <= 1e-5 * float(tot) ** 2
    for t in (0, 1, len(numels) // 2, len(numels) - 1):
        ref = float(torch.linalg.vector_norm(g[t], dtype=torch.float64))
        assert abs(float(per[t]) - ref) <= 1e-5 * ref
    # power-of-two scaling is exact: bf16 -> fp32 * 2^-10, then fp32 -> bf16 * 2^10 returns the gradient bits.
    # (done on the largest tensor to bound the extra memory: 525 M elements)
    big = max(range(len(numels)), key=lambda i: numels[i])
    tmp32 = torch.empty(numels[big], dtype=F32, device="cuda")
    back = torch.empty(numels[big], dtype=BF16, device="cuda")
    amp_C.multi_tensor_scale(8192, flag, [[g[big]], [tmp32]], 2.0 ** -10)
    amp_C.multi_tensor_scale(8192, flag, [[tmp32], [back]], 2.0 ** 10)
    assert flag.item() == 0 and torch.equal(back.view(torch.int16), g[big].view(torch.int16))
    # one inf at the very last element of the 8 G-element set is seen
    last = g[-1]
    keep = last[-1].clone()
    last[-1] = float("inf")
    amp_C.multi_tensor_l2norm(65536, flag, [g], False)
    assert flag.item() == 1
    last[-1] = keep


# ------------------------------------------------------------------------------------------------ C2: GPT-2-small amp step
def test_gpt2_small_amp_step_full_oracle(amp_C, oracle):
    """C2 in full (124 439 808 params, 148 tensors): fp16 gradients carrying loss scale 65536, fp32 masters, fp16 model
    copy; inf/norm check then the fused mixed-precision Adam -- the oracle runs over EVERY element.  Overflow variant:
    inf at element 0 of tensor 0 and nan at the last element of the last tensor => flag 1 and nothing moves."""
    numels = workloads.numels(workloads.gpt2_small())
    assert sum(numels) == 124439808 and len(numels) == 148
    P, nt = sum(numels), len(numels)
    gen = torch.Generator().manual_seed(9876)
    hp_, hm_, hv_ = (torch.empty(P).normal_(0, s, generator=gen) for s in (0.02, 1e-3, 1e-3))
    hv_.square_()
    hg_ = (torch.empty(P).normal_(0, 1e-3, generator=gen) * 65536).to(F16)
    host = [_carve(x, numels) for x in (hp_, hm_, hv_, hg_, torch.zeros(P, dtype=F16))]
    dev = [_carve(x.cuda(), numels) for x in (hp_, hm_, hv_, hg_, torch.zeros(P, dtype=F16))]
    T = lambda val, dt=F32: torch.full((nt,), val, dtype=dt, device="cuda")
    hp = (T(0.9), T(0.999), T(1, torch.int32), T(1e-8), T(0.01))
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    gn, _ = amp_C.multi_tensor_l2norm(65536, flag, [dev[3]], False)
    amp_C.multi_tensor_fused_adam(8192, flag, dev, *hp, 5e-4, 65536.0, 3, 1, True)
    assert flag.item() == 0
    ref = float(torch.linalg.vector_norm(hg_, dtype=torch.float64))
    assert abs(float(gn) - ref) <= 1e-5 * ref
    oracle.multi_tensor_fused_adam(8192, torch.zeros(1, dtype=torch.int32), host, [0.9] * nt, [0.999] * nt, [1] * nt,
                                   [1e-8] * nt, [0.01] * nt, 5e-4, 65536.0, 3, 1)
    for d in (0, 1, 2, 4):
        got = torch.cat([t.reshape(-1) for t in dev[d]]).cpu()
        want = torch.cat([t.reshape(-1) for t in host[d]])
        assert bits_equal(got, want), f"list {d}: {n_mismatch(got, want)} of {P} elements differ"
    # overflow variant
    dev[3][0][0] = float("inf")
    dev[3][-1][-1] = float("nan")
    before = [_bit_checksum(torch.cat([t.reshape(-1) for t in dev[d]])) for d in (0, 1, 2, 4)]
    flag.zero_()
    amp_C.multi_tensor_l2norm(65536, flag, [dev[3]], False)
    assert flag.item() == 1
    amp_C.multi_tensor_fused_adam(8192, flag, dev, *hp, 5e-4, 65536.0, 4, 1, True)
    assert [_bit_checksum(torch.cat([t.reshape(-1) for t in dev[d]])) for d in (0, 1, 2, 4)] == before


# ------------------------------------------------------------------------------------------------ C3: 50 000 tensors
def test_c3_50k_tensors_sgd_and_scale(amp_C, oracle):
    """C3 in full (50 000 tensors, ~1.07 G elements): ONE launch each; every 97th tensor (all 13 sizes of the cycle occur)
    is checked bit for bit against the oracle; an integer checksum shows the unsampled rest moved too."""
    _need_gb(24)
    shapes = workloads.c3_many_small()
    numels = workloads.numels(shapes)
    assert len(numels) == 50000
    P = sum(numels)
    gen = torch.Generator(device="cuda").manual_seed(5)
    fg, fw, fm = (torch.empty(P, dtype=F32, device="cuda").normal_(0, s, generator=gen) for s in (1e-2, 0.05, 1e-3))
    g, w, mom = (_carve(x, numels) for x in (fg, fw, fm))
    idx = list(range(0, 50000, 97))
    host = [[l[i].cpu().clone() for i in idx] for l in (g, w, mom)]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    args = (1e-4, 0.9, 0.0, 0.25, True, False, False, 1.0)
    c0 = _bit_checksum(fw)
    launches = amp_C.launch_count()
    amp_C.multi_tensor_sgd(8192, flag, [g, w, mom], *args)
    assert amp_C.launch_count() - launches == 1
    assert _bit_checksum(fw) != c0
    oracle.multi_tensor_sgd(8192, torch.zeros(1, dtype=torch.int32), host, *args)
    for d, l in ((1, w), (2, mom)):
        for j, i in enumerate(idx):
            assert bits_equal(l[i], host[d][j]), f"list {d} tensor {i} (n={numels[i]})"
    # multi_tensor_scale fp32 -> fp16 -> fp32 by powers of two over all 50 000 tensors: exact on fp16-representable data
    f16 = torch.empty(P, dtype=F16, device="cuda")
    h = _carve(f16, numels)
    fq = fw.to(F16).float()                       # fp16-representable fp32 values
    q = _carve(fq, numels)
    out = torch.empty(P, dtype=F32, device="cuda")
    o = _carve(out, numels)
    amp_C.multi_tensor_scale(8192, flag, [q, h], 4.0)
    amp_C.multi_tensor_scale(8192, flag, [h, o], 0.25)
    assert flag.item() == 0
    big = fq.abs() < 8000                          # x4 must stay finite in fp16
    assert bool(big.all()) and torch.equal(out, fq)


# ------------------------------------------------------------------------------------------------ C4: BERT-large LAMB
@pytest.mark.parametrize("use_nvlamb", [False, True])
def test_bert_large_lamb_full_oracle(amp_C, oracle, use_nvlamb):
    """C4 in full (~336 M params, fp32): global-norm pass + multi_tensor_lamb; the oracle runs over every element.
    m, v: bit-exact (no norm involved); per-tensor ||p||, ||u|| and the global ||g||: 1e-5 relative; p: 1e-6."""
    _need_gb(16)
    numels = workloads.numels(workloads.bert_large())
    P, nt = sum(numels), len(numels)
    gen = torch.Generator().manual_seed(9876)
    hg_, hp_, hm_, hv_ = (torch.empty(P).normal_(0, s, generator=gen) for s in (1e-3, 0.02, 1e-4, 1e-4))
    hv_.square_()
    host = [_carve(x, numels) for x in (hg_, hp_, hm_, hv_)]
    dev = [_carve(x.cuda(), numels) for x in (hg_, hp_, hm_, hv_)]
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    gn, _ = amp_C.multi_tensor_l2norm(65536, flag, [dev[0]], False)
    ref_gn = float(torch.linalg.vector_norm(hg_, dtype=torch.float64))
    assert abs(float(gn) - ref_gn) <= 1e-5 * ref_gn
    args = (5e-4, 0.9, 0.999, 1e-6, 3, 1, 0.01, 1, 1)
    launches = amp_C.launch_count()
    amp_C.multi_tensor_lamb(8192, flag, dev, *args, gn, 1.0, use_nvlamb)
    assert amp_C.launch_count() - launches == 2          # stage 1 (+ both per-tensor norms, finalized in-launch), stage 2
    # same clip decision on both sides: the norm is well above max_grad_norm
    assert ref_gn > 1.0 and float(gn) > 1.0
    pn, un = oracle.multi_tensor_lamb(8192, torch.zeros(1, dtype=torch.int32), host, *args, gn.cpu(), 1.0, use_nvlamb,
                                      return_norms=True)
    for d in (2, 3):
        got = torch.cat([t.reshape(-1) for t in dev[d]]).cpu()
        want = torch.cat([t.reshape(-1) for t in host[d]])
        assert bits_equal(got, want), f"list {d}: {n_mismatch(got, want)} of {P} elements differ"
    got = torch.cat([t.reshape(-1) for t in dev[1]]).cpu()
    want = torch.cat([t.reshape(-1) for t in host[1]])
    assert torch.allclose(got, want, rtol=1e-6, atol=1e-9)
    # the update term left in g (stage 1 output): bit-exact as well
    got = torch.cat([t.reshape(-1) for t in dev[0]]).cpu()
    want = torch.cat([t.reshape(-1) for t in host[0]])
    assert bits_equal(got, want)
    # per-tensor update norms the oracle used (host[0] now holds u) vs double precision: 1e-5
    for i in (0, 5, nt - 1):
        r = float(torch.linalg.vector_norm(host[0][i], dtype=torch.float64))
        assert abs(float(un[i]) - r) <= 1e-5 * r


# ------------------------------------------------------------------------------------------------ C1: 1 M-param MLP
@pytest.mark.parametrize("adam_w_mode,wd", [(False, 0.0), (True, 0.01)])
def test_c1_mlp_fused_adam_vs_torch_optim_cpu(adam_w_mode, wd):
    """C1: the ~1 M-parameter fp32 MLP (986 634 params, 6 tensors) trained for 7 steps with FusedAdam on the B200 against
    torch.optim.Adam / AdamW on the CPU fed the very same gradient bits (the reference's own CPU-runnable oracle,
    tests/L0/run_optimizers/test_fused_optimizer.py:64-88).  Tolerance 1e-6 relative (north star), and the reference
    test's 1e-3 absolute."""
    from beforeholiday_b200.optimizers import FusedAdam

    torch.manual_seed(9876)
    mk = lambda: torch.nn.Sequential(torch.nn.Linear(1024, 768), torch.nn.ReLU(), torch.nn.Linear(768, 256), torch.nn.ReLU(),
                                     torch.nn.Linear(256, 10))
    cpu_model = mk()
    gpu_model = mk().cuda()
    gpu_model.load_state_dict(cpu_model.state_dict())
    assert sum(p.numel() for p in cpu_model.parameters()) == 986634
    opts = dict(lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    ref = (torch.optim.AdamW if adam_w_mode else torch.optim.Adam)(cpu_model.parameters(), **opts)
    tst = FusedAdam(gpu_model.parameters(), adam_w_mode=adam_w_mode, **opts)
    for it in range(7):
        x = torch.randn(64, 1024, device="cuda")
        y = torch.randint(0, 10, (64,), device="cuda")
        tst.zero_grad()
        torch.nn.functional.cross_entropy(gpu_model(x), y).backward()
        for pc, pg in zip(cpu_model.parameters(), gpu_model.parameters()):
            pc.grad = pg.grad.detach().cpu().clone()
        tst.step()
        ref.step()
    for pc, pg in zip(cpu_model.parameters(), gpu_model.parameters()):
        a, b = pg.detach().cpu(), pc.detach()
        assert float((a - b).abs().max()) <= 1e-3
        assert torch.allclose(a, b, rtol=1e-6, atol=2e-7)
