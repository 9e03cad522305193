"""Single-process replay of sharded LAMB with the CPU oracle on WHOLE parameters (test infrastructure): what every
rank's parameters must equal after DistributedFusedLAMB steps.  Norms are taken in double over the whole parameter, so
the sharded product agrees to summation-order rounding (norms 1e-5 relative => parameters ~1e-6 relative)."""
import torch


def make_grad(seed, n, dtype, scale):
    """Small dyadic values: their sum over the ranks and the 1/world average are exact in every dtype."""
    g = torch.Generator().manual_seed(seed)
    return (torch.randint(-16, 17, (n,), generator=g).float() / 64 * scale).to(dtype)


def replay(O, sizes, init, props, world, steps, dtype, lr, global_scale, max_grad_norm, mode, use_nvlamb, grad_seed,
           overflow_at=None, e5m2=False, clip_after_ar=True):
    """props[i] = (weight_decay, bias_correction, beta1, beta2, eps).  Returns (masters, model-dtype copies, step).
    e5m2: the model parameters are the e5m2-packed masters (reference e5m2_allgather).  clip_after_ar=False: every rank
    clips its LOCAL gradients by their own norm and pre-divides by the world size, the reduction is a plain sum and
    stage 1 does not clip (reference distributed_fused_lamb.py:629-646, :697)."""
    k = len(sizes)
    norm_dt = torch.uint8 if e5m2 else dtype
    p32 = [x.to(dtype).float() for x in init]
    m, v, u = ([torch.zeros(n) for n in sizes] for _ in range(3))
    pc = [x.to(dtype).clone() for x in init]
    step = 0
    for it in range(steps):
        gs = [torch.zeros(n) for n in sizes]
        for r in range(world):
            local = [make_grad(grad_seed(it, r, i), n, dtype, global_scale) for i, n in enumerate(sizes)]
            if clip_after_ar:
                for i in range(k):
                    gs[i] += local[i].float() / world
            else:
                norm = torch.cat([x.double() for x in local]).norm().float()
                coeff = torch.tensor(global_scale * max_grad_norm, dtype=torch.float32) / (1e-6 + norm)
                coeff = coeff if float(coeff) <= 1 else torch.ones(())
                scale = (coeff.half() / world).to(dtype)
                for i in range(k):
                    gs[i] += (local[i] * scale).float()          # dtype x dtype product, rounded to dtype, summed
        gs = [x.to(dtype) for x in gs]
        if overflow_at == it:
            continue                                           # the step is skipped: nothing moves, step stays
        gnorm = torch.cat([x.double() for x in gs]).norm().float()
        step += 1
        flag = torch.zeros(1, dtype=torch.int32)
        pn = [float(O.convert(O.convert(x, norm_dt), torch.float32).double().norm()) for x in p32]
        O.multi_tensor_lamb_compute_update_term(8192, flag, [gs, p32, m, v, u], [q[2] for q in props], [q[3] for q in props],
                                                [0.0] * k, [q[1] for q in props], step, [q[4] for q in props], mode,
                                                [q[0] for q in props], global_scale, gnorm,
                                                max_grad_norm if clip_after_ar else 0.0)
        un = [float(x.double().norm()) for x in u]
        copies = [torch.zeros(n, dtype=torch.uint8) for n in sizes] if e5m2 else pc
        O.multi_tensor_lamb_update_weights(8192, flag, [u, p32, copies], pn, un, list(range(k)), lr, [q[0] for q in props],
                                           gnorm, use_nvlamb)
        if e5m2:
            pc = [O.convert(O.convert(c, torch.float32), dtype) for c in copies]
    return p32, pc, step
