"""The product's coalition sampler against the oracle's restatement (CPU only): same seed,
same random-stream consumption -> identical masks and weights."""

import numpy as np
import pytest

from fraud_detection_b200 import sampler
from oracle import kernel_shap as ks


@pytest.mark.parametrize("M,nsamples", [(30, 2048), (30, 4096), (30, "auto"), (30, 100), (12, 500), (7, "auto"), (4, "auto"), (2, "auto")])
def test_same_masks_as_oracle(M, nsamples):
    bits, kw = sampler.sample_coalitions(M, nsamples, np.random.RandomState(11))
    Z, kw_ref, info = ks.sample_coalitions(M, nsamples, np.random.RandomState(11))
    assert np.array_equal(bits, sampler.masks_to_bits(Z))
    assert np.array_equal(kw, kw_ref)
    assert np.array_equal(sampler.bits_to_masks(bits, M), Z)
    assert len(bits) <= info["nsamples"]


def test_global_state_is_consumed_like_the_reference():
    np.random.seed(3)
    a1 = sampler.sample_coalitions(30, 2048)
    a2 = sampler.sample_coalitions(30, 2048)          # a second row gets different masks
    np.random.seed(3)
    b1 = ks.sample_coalitions(30, 2048)
    b2 = ks.sample_coalitions(30, 2048)
    assert np.array_equal(a1[0], sampler.masks_to_bits(b1[0])) and np.array_equal(a2[0], sampler.masks_to_bits(b2[0]))
    assert not np.array_equal(a1[0], a2[0])


def test_bit_roundtrip_and_validation():
    Z = (np.random.RandomState(0).rand(50, 30) > 0.5).astype(float)
    assert np.array_equal(sampler.bits_to_masks(sampler.masks_to_bits(Z)), Z)
    with pytest.raises(ValueError):
        sampler.masks_to_bits(np.zeros((3, 31)))


def test_partially_varying_layout_and_bit_spreading():
    """Host logic behind fdb_kshap_plan_create_varying: the last varying feature moves to index 29 (the one the
    constraint eliminates), coalition bit j lands on the j-th varying feature's position."""
    from fraud_detection_b200.explainers import B200KernelExplainer as KE
    from fraud_detection_b200 import sampler
    for v in ([0, 3, 4, 9, 17, 22, 26], [5, 29], [1, 2, 28], list(range(29)), [0, 29]):
        perm, pos, mask = KE._varying_layout(v)
        assert sorted(perm.tolist()) == list(range(30))
        assert perm[29] == v[-1] and perm[v[-1]] == 29                    # a swap, nothing else moves
        assert all(perm[j] == j for j in range(30) if j not in (29, v[-1]))
        assert pos[-1] == 29 and pos[:-1].tolist() == list(v[:-1])
        assert mask == sum(1 << int(p) for p in pos) and (mask >> 29) & 1
        # the permuted column at bit pos[j] is the j-th varying feature of the original row
        assert [int(perm[p]) for p in pos] == list(v)
        M = len(v)
        bits, kw = sampler.sample_coalitions(M, 64, np.random.RandomState(M))
        Z = sampler.bits_to_masks(bits, M)
        wide = KE._spread_bits(bits, pos)
        assert np.all((wide & ~np.uint32(mask)) == 0)
        for j, pz in enumerate(pos):
            assert np.array_equal((wide >> np.uint32(pz)) & 1, Z[:, j].astype(np.uint32))


def test_host_near_gpu_is_a_noop_without_a_device():
    import os
    from fraud_detection_b200 import host_near_gpu
    before = os.sched_getaffinity(0)
    with host_near_gpu(0) as near:
        assert near.cpus is None or set(near.cpus) <= before
    assert os.sched_getaffinity(0) == before
