"""Property tests (hypothesis) of the sharded optimizers' flat layout: for arbitrary parameter sizes, world sizes,
element widths and bucket caps, every flat element is owned by exactly one rank, every parameter element lies in
exactly one fragment, fragments stay inside their sub-shard and map to the right state offset, sub-shards start on
256-element boundaries, and the fragments of one parameter are consecutive on a rank (what DistributedFusedLAMB's
per-parameter norm runs rely on)."""
from hypothesis import given, settings
from hypothesis import strategies as st

from beforeholiday_b200.contrib.optimizers import FlatLayout


@settings(max_examples=150, deadline=None)
@given(numels=st.lists(st.integers(min_value=1, max_value=5000), min_size=1, max_size=24),
       world=st.sampled_from([1, 2, 3, 4, 8]), esz=st.sampled_from([2, 4]),
       cap=st.integers(min_value=1, max_value=20000))
def test_flat_layout_invariants(numels, world, esz, cap):
    layouts = [FlatLayout(numels, world, r, elem_bytes=esz, bucket_cap_elems=cap) for r in range(world)]
    L0 = layouts[0]
    align = 128 // esz
    assert all(o % align == 0 for o in L0.offsets)                        # 128-byte aligned parameters
    assert all(b >= a + n for a, n, b in zip(L0.offsets, numels, L0.offsets[1:] + [L0.total]))   # no overlap
    assert L0.padded == L0.nbuckets * L0.bucket_elems >= L0.total
    assert L0.bucket_elems == world * L0.sub_shard and L0.sub_shard % 256 == 0
    assert L0.state_elems == L0.nbuckets * L0.sub_shard
    owner = [0] * L0.padded
    covered = {}
    for r, L in enumerate(layouts):
        assert (L.offsets, L.padded, L.nbuckets, L.sub_shard) == (L0.offsets, L0.padded, L0.nbuckets, L0.sub_shard)
        for b in range(L.nbuckets):
            lo, hi = L.owned_range(b)
            blo, bhi = L.bucket_range(b)
            assert blo <= lo < hi <= bhi and lo % 256 == 0
            for i in range(lo, hi):
                owner[i] += 1
        frags = L.fragments()
        last_seen = {}
        for k, f in enumerate(frags):
            n = f.flat_hi - f.flat_lo
            assert n > 0 and L.offsets[f.param] + f.param_lo == f.flat_lo and f.param_lo + n <= numels[f.param]
            olo, ohi = L.owned_range(f.bucket)
            assert olo <= f.flat_lo and f.flat_hi <= ohi
            assert f.state_lo == f.bucket * L.sub_shard + (f.flat_lo - olo)
            if f.param in last_seen:
                assert last_seen[f.param] == k - 1, "fragments of one parameter must be consecutive on a rank"
            last_seen[f.param] = k
            for i in range(f.param_lo, f.param_lo + n):
                covered[(f.param, i)] = covered.get((f.param, i), 0) + 1
        # fragments are listed in flat order
        assert [f.flat_lo for f in frags] == sorted(f.flat_lo for f in frags)
    assert all(c == 1 for c in owner)
    assert len(covered) == sum(numels) and all(c == 1 for c in covered.values())
