"""Synthetic parameter sets for BASELINE.json's configs (shapes only).

Used by bench.py and the tests.  Nothing here touches a GPU.
"""
from __future__ import annotations

from math import prod


def c1_mlp():
    """C1: ~1M-param fp32 MLP (986 634 params, 6 tensors)."""
    return [(768, 1024), (768,), (256, 768), (256,), (10, 256), (10,)]


def gpt2_small():
    """C2: GPT-2 small, 124 439 808 params, 148 tensors."""
    s = [(50257, 768), (1024, 768)]
    for _ in range(12):
        s += [(768,), (768,), (768, 2304), (2304,), (768, 768), (768,), (768,), (768,), (768, 3072), (3072,),
              (3072, 768), (768,)]
    s += [(768,), (768,)]
    return s


def c3_many_small(n_tensors=50000):
    """C3: tensor-count stress, sizes cycling through BN/LN vectors, biases, 1x1 and 3x3 conv kernels."""
    cyc = [64, 128, 256, 512, 1024, 2048, 768, 3072, 4096, 16384, 36864, 65536, 147456]
    return [(cyc[i % len(cyc)],) for i in range(n_tensors)]


def bert_large():
    """C4: BERT-large (~336M params, ~398 tensors)."""
    h, i, v = 1024, 4096, 30522
    s = [(v, h), (512, h), (2, h), (h,), (h,)]
    for _ in range(24):
        s += [(h, h), (h,)] * 4 + [(h,), (h,)] + [(i, h), (i,), (h, i), (h,)] + [(h,), (h,)]
    s += [(h, h), (h,)]                      # pooler
    s += [(h, h), (h,), (h,), (h,), (v,)]    # MLM head: dense, LN, decoder bias
    s += [(2, h), (2,)]                      # NSP
    return s


def llama3_8b():
    """C5: Llama-3-8B, 8 030 261 248 params, 291 tensors."""
    h, kv, ff, v = 4096, 1024, 14336, 128256
    s = [(v, h)]
    for _ in range(32):
        s += [(h, h), (kv, h), (kv, h), (h, h), (ff, h), (ff, h), (h, ff), (h,), (h,)]
    s += [(h,), (v, h)]
    return s


WORKLOADS = {
    "c1_mlp": c1_mlp,
    "gpt2_small": gpt2_small,
    "c3_many_small": c3_many_small,
    "bert_large": bert_large,
    "llama3_8b": llama3_8b,
}


def numels(shapes):
    return [prod(s) for s in shapes]


def total(shapes):
    return sum(numels(shapes))
