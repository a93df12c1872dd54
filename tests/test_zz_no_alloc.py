"""bm_cosine_distance_noalloc: the allocation-free variant of bm_cosine_distance (static
__device__ scratch inside the library, see include/boxmot_b200.h).  Collected last on purpose:
a failure here must not hide any other test.

The check is against the EXISTING entry point on the same inputs, on two calls in a row that
share the static scratch (the second with a different, smaller shape, so stale contents of the
first call lie behind and under it)."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _inputs(na, nb, E, seed):
    rng = np.random.default_rng(seed)
    a = rng.normal(0, 1, (na, E)).astype(np.float32)
    b = rng.normal(0, 1, (nb, E)).astype(np.float32)
    b[: min(na, nb)] += 3 * a[: min(na, nb)]
    return torch.as_tensor(a).cuda(), torch.as_tensor(b).cuda()


@pytest.mark.parametrize("mode", [0, 1])
def test_noalloc_cosine_matches_existing_entry_point_twice_in_a_row(mode):
    from boxmot_b200 import _lib
    L = _lib.lib()
    # tolerance of test_cosine_distance_matches_scipy (tests/test_gpu_units.py); the two entry
    # points launch the same kernels, so the observed difference is 0
    atol = 2e-6 if mode else 2e-5
    for call, (na, nb, E) in enumerate([(300, 260, 96), (37, 130, 100)]):
        d_a, d_b = _inputs(na, nb, E, 100 + call)
        want = torch.full((na, nb), -7.0, dtype=torch.float32, device="cuda")
        work = torch.zeros((2 * (na + nb),), dtype=torch.int32, device="cuda")
        _lib.check(L.bm_cosine_distance(_p(d_a), na, _p(d_b), nb, E, _p(want), _p(work), mode, None), "cosine")
        got = torch.full((na, nb), -9.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        _lib.check(L.bm_cosine_distance_noalloc(_p(d_a), na, _p(d_b), nb, E, _p(got), mode, None), "cosine noalloc")
        torch.cuda.synchronize()
        w, g = want.cpu().numpy(), got.cpu().numpy()
        assert np.isfinite(g).all() and g.min() >= 0.0
        np.testing.assert_allclose(g, w, rtol=0, atol=atol, err_msg="call %d" % call)


def test_noalloc_cosine_refuses_sizes_beyond_the_static_scratch():
    from boxmot_b200 import _lib
    L = _lib.lib()
    cap = 16384                      # BM_COSINE_NOALLOC_MAX_ROWS (include/boxmot_b200.h)
    d_a, d_b = _inputs(8, 8, 32, 7)
    out = torch.zeros((8, 8), dtype=torch.float32, device="cuda")
    rc = L.bm_cosine_distance_noalloc(_p(d_a), cap, _p(d_b), 1, 32, _p(out), 1, None)
    assert rc == -4
    assert b"BM_COSINE_NOALLOC_MAX_ROWS" in L.bm_last_error()
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0     # nothing was launched
