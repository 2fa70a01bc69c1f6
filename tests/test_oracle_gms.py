"""CPU: the GMS restatement (oracle/gms_oracle.c) against the committed golden masks -- produced by the reference's own
MatchGMS.cpp compiled from /root/reference (oracle/gen_golden_gms.py) -- and, where that build is present, against it live."""
import os

import numpy as np
import pytest

import datagen_gms
from oracle import gms

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "gms_golden.npz"))


@pytest.mark.parametrize("name", sorted(datagen_gms.CASES))
def test_oracle_equals_reference_golden(name):
    kw = datagen_gms.CASES[name]
    k1, k2, m, sz = datagen_gms.scene(**kw["scene"])
    n, mask = gms.filter_matches(k1, sz, k2, sz, m, kw["use_scale"], kw["use_rotation"])
    assert n == int(G[name + "_n"]) and np.array_equal(mask, G[name + "_mask"])
    assert n == int(mask.sum())


def test_oracle_equals_reference_build_live():
    gms.build(ref=True)
    if gms.ref_lib() is None:
        pytest.skip("no /root/reference here: oracle/_ref/libgms_ref.so not built")
    for seed in range(20, 32):
        k1, k2, m, sz = datagen_gms.scene(n=500 + 331 * seed, seed=seed, frac=0.3 + 0.05 * (seed % 8), rot=0.1 * (seed % 3), sc=1.0 + 0.1 * (seed % 4))
        for us, ur in ((0, 0), (seed % 2, (seed // 2) % 2)):
            a = gms.filter_matches(k1, sz, k2, sz, m, us, ur); b = gms.ref_filter_matches(k1, sz, k2, sz, m, us, ur)
            assert a[0] == b[0] and np.array_equal(a[1], b[1]), (seed, us, ur)


def test_empty_and_inlier_free():
    k1, k2, m, sz = datagen_gms.scene(n=100, seed=1, frac=0.0)
    n, mask = gms.filter_matches(k1, sz, k2, sz, m)
    assert n == 0 and not mask.any()
    n, mask = gms.filter_matches(k1, sz, k2, sz, m[:0])
    assert n == 0 and len(mask) == 0
