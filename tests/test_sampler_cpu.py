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
