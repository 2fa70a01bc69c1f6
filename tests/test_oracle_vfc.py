"""CPU: the VFC checker -- the reference's own vfc.cpp / vfcMatches.cpp (oracle/_ref/libvfc_ref.so) separates a smooth motion field
from random matches on seeded scenes, and returns the codes vfcMatches.cpp documents."""
import numpy as np
import pytest

import datagen_vfc
from oracle import vfc as ovfc


def _built():
    ovfc.build()
    if ovfc.ref_lib() is None:
        pytest.skip("oracle/_ref/libvfc_ref.so not built (needs /root/reference)")


@pytest.mark.parametrize("n,outl,seed", [(3000, 0.4, 1), (500, 0.2, 2), (30, 0.1, 4)])
def test_reference_vfc_separates_the_field(n, outl, seed):
    _built()
    k1, k2, mt, inl = datagen_vfc.scene(n, outl, seed=seed)
    code, keep = ovfc.filter_with_vfc(k1, k2, mt, seed=1)
    keep = keep.astype(bool)
    assert code == 0 and (keep & inl).sum() >= 0.98 * inl.sum() and (keep & ~inl).sum() <= 0.02 * n + 1


def test_reference_vfc_codes():
    _built()
    k1, k2, mt, inl = datagen_vfc.scene(40, 0.1, seed=9)
    assert ovfc.filter_with_vfc(k1, k2, mt[:4], seed=1)[0] == -1                 # fewer than MIN_POINT_NUMBER
    a = ovfc.filter_with_vfc(k1, k2, mt, seed=5); b = ovfc.filter_with_vfc(k1, k2, mt, seed=5)
    assert a[0] == b[0] and np.array_equal(a[1], b[1])                          # the control points come from rand()
