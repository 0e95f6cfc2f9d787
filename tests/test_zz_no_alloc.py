"""glia_b200_rf_score_noalloc (include/glia_b200.h): the allocation-free RF scoring entry point whose vote-table scratch
is a static __device__ array.  No counterpart in the reference.  Checked
through the C ABI against the existing entry point glia_b200_rf_score on two calls in a row (the second smaller and on
other rows: stale counters of the first call must not leak), plus the size limit.  RF probabilities are votes / ntree,
so the two paths must agree exactly (tolerance 0, as in the other RF tests).

This file sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

from glia_b200 import api, synth

pytestmark = pytest.mark.gpu


def _score_both(forest, X, label):
    import torch
    L = api.lib()
    n, d = X.shape
    dX = torch.from_numpy(X).cuda()
    want = torch.full((n,), -1.0, dtype=torch.float64, device="cuda")
    got = torch.full((n,), -2.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    st = L.glia_b200_rf_score(forest.h, C.c_void_p(dX.data_ptr()), n, d, label, None, C.c_void_p(want.data_ptr()), 1, None)
    assert st == 0
    err = C.create_string_buffer(256)
    st = L.glia_b200_rf_score_noalloc(forest.h, C.c_void_p(dX.data_ptr()), n, d, label, C.c_void_p(got.data_ptr()), None, err, 256)
    assert st == 0, err.value
    torch.cuda.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


@pytest.mark.parametrize("nclass,labels,label", [(2, (-1, 1), -1), (3, (-1, 1, 2), 1)])
def test_noalloc_matches_rf_score_on_two_calls_in_a_row(nclass, labels, label):
    d = 37
    rng = np.random.default_rng(5)
    pool = rng.random((2000, d))
    f = synth.random_forest(61, d, max_depth=10, seed=3, nclass=nclass, orig_labels=labels, thr_pool=pool)
    forest = api.Forest.from_arrays(f, d)
    X1 = rng.random((5000, d))
    X1[::7] = pool[rng.integers(0, len(pool), size=len(X1[::7]))]   # rows sitting exactly on thresholds (<= goes left)
    X1[3, 5] = np.nan; X1[4, 6] = np.inf; X1[5, 7] = -np.inf
    want, got = _score_both(forest, X1, label)
    assert np.array_equal(got, want)
    assert 0.0 < want.mean() < 1.0
    X2 = rng.random((1234, d))                                      # second call: fewer, different rows
    want2, got2 = _score_both(forest, X2, label)
    assert np.array_equal(got2, want2)
    want3, got3 = _score_both(forest, X1, label)                    # and the first batch again
    assert np.array_equal(got3, want) and np.array_equal(want3, want)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    L = api.lib()
    d = 4
    f = synth.random_forest(5, d, max_depth=3, seed=1)
    forest = api.Forest.from_arrays(f, d)
    cap = L.glia_b200_rf_score_noalloc_max_votes()
    assert cap == 1 << 20
    n = cap // forest.nclass + 1
    dX = torch.zeros((n, d), dtype=torch.float64, device="cuda")
    out = torch.zeros(n, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(256)
    st = L.glia_b200_rf_score_noalloc(forest.h, C.c_void_p(dX.data_ptr()), n, d, -1, C.c_void_p(out.data_ptr()), None, err, 256)
    assert st == 3 and b"exceed the static scratch" in err.value   # GLIA_E_CAPACITY
    st = L.glia_b200_rf_score_noalloc(forest.h, C.c_void_p(dX.data_ptr()), n - 1, d, -1, C.c_void_p(out.data_ptr()), None, err, 256)
    assert st == 0, err.value                                       # exactly the maximum is accepted
    torch.cuda.synchronize()
    st = L.glia_b200_rf_score_noalloc(forest.h, C.c_void_p(dX.data_ptr()), 8, d, 12345, C.c_void_p(out.data_ptr()), None, err, 256)
    assert st == 1 and err.value                                     # unknown class label: GLIA_E_ARG with a message
