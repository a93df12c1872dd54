"""Function-level GPU parity: unit kernels called through the C ABI vs the oracle."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev(a, dtype=None):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda()


def _lap_batch(costs, limit=np.inf, entry="bm_lap_solve_batch"):
    from boxmot_b200 import _lib
    B = len(costs)
    nr = np.array([c.shape[0] for c in costs], np.int32)
    nc = np.array([c.shape[1] for c in costs], np.int32)
    R, C = max(int(nr.max()), 1), max(int(nc.max()), 1)
    buf = np.zeros((B, R, C))
    for b, c in enumerate(costs):
        buf[b, : c.shape[0], : c.shape[1]] = c
    d_c, d_nr, d_nc = _dev(buf), _dev(nr), _dev(nc)
    d_x = torch.full((B, R), -7, dtype=torch.int32, device="cuda")
    d_y = torch.full((B, C), -7, dtype=torch.int32, device="cuda")
    lim = float(limit) if limit < np.inf else float("inf")
    _lib.check(getattr(_lib.lib(), entry)(
        ctypes.c_void_p(d_c.data_ptr()), B, R, C, ctypes.c_void_p(d_nr.data_ptr()),
        ctypes.c_void_p(d_nc.data_ptr()), lim, ctypes.c_void_p(d_x.data_ptr()),
        ctypes.c_void_p(d_y.data_ptr()), None), "bm_lap_solve_batch")
    torch.cuda.synchronize()
    return d_x.cpu().numpy(), d_y.cpu().numpy()


def _random_costs(rng, n, kind):
    out = []
    for it in range(n):
        nr, nc = rng.integers(1, 90, 2)
        c = rng.random((nr, nc))
        k = kind if kind >= 0 else it % 5
        if k == 1:
            c = np.round(c * 4) / 4                      # heavy exact ties
        elif k == 2:
            c = -np.where(rng.random((nr, nc)) < 0.9, 0.0, c)   # IoU-like: mostly exact zeros
        elif k == 3:
            c = -(np.where(rng.random((nr, nc)) < 0.85, 0.0, c) + 0.02 * np.round(rng.normal(size=(nr, nc))))
        elif k == 4:
            c = np.zeros((nr, nc))                       # fully degenerate
        out.append(c)
    return out


@pytest.mark.parametrize("limit", [np.inf, 0.9, 0.3])
def test_lap_batch_equals_lapjv_restatement_including_ties(limit):
    from oracle import lap
    rng = np.random.default_rng(5)
    costs = _random_costs(rng, 400, -1)
    x, y = _lap_batch(costs, limit)
    for b, c in enumerate(costs):
        _, xo, yo = lap.lapjv(c, extend_cost=True, cost_limit=limit)
        assert np.array_equal(x[b, : c.shape[0]], xo), "problem %d rows" % b
        assert np.array_equal(y[b, : c.shape[1]], yo), "problem %d cols" % b
        assert (x[b, c.shape[0]:] == -1).all() and (y[b, c.shape[1]:] == -1).all()


def test_lap_batch_large_tracking_shaped():
    """C2/C5-shaped problems: D x T IoU-like costs, n_ext up to ~700."""
    from oracle import lap
    rng = np.random.default_rng(9)
    costs = []
    for nr, nc in [(41, 145), (100, 100), (200, 260), (300, 400), (1, 300), (300, 1)]:
        c = -np.where(rng.random((nr, nc)) < 0.95, 0.0, rng.random((nr, nc)))
        c += -0.05 * (rng.random((nr, nc)) - 0.5) * (rng.random((1, nc)) < 0.6)
        costs.append(c)
    x, y = _lap_batch(costs)
    for b, c in enumerate(costs):
        _, xo, yo = lap.lapjv(c, extend_cost=True)
        assert np.array_equal(x[b, : c.shape[0]], xo)
        assert np.array_equal(y[b, : c.shape[1]], yo)


@pytest.mark.parametrize("limit", [np.inf, 0.9, 0.3, -0.2])
def test_lap_rect_equals_lapjv_on_unique_optima(limit):
    """The engine's fast solver (real rows + implicit null option) returns lapjv's x / y
    whenever the optimum is unique (continuous random costs), for every cost_limit mode."""
    from oracle import lap
    rng = np.random.default_rng(11)
    costs = []
    for it in range(300):
        nr, nc = rng.integers(1, 120, 2)
        c = rng.random((nr, nc)) * 2.0 - 1.2
        if it % 3 == 0:      # tracking-shaped: few strong pairs on a weak dense background
            c = -0.05 * rng.random((nr, nc))
            k = min(nr, nc)
            c[rng.permutation(nr)[:k], rng.permutation(nc)[:k]] -= rng.uniform(0.3, 1.0, k)
        costs.append(c)
    x, y = _lap_batch(costs, limit, entry="bm_lap_rect_batch")
    for b, c in enumerate(costs):
        _, xo, yo = lap.lapjv(c, extend_cost=True, cost_limit=limit)
        assert np.array_equal(x[b, : c.shape[0]], xo), "problem %d rows" % b
        assert np.array_equal(y[b, : c.shape[1]], yo), "problem %d cols" % b
        assert (x[b, c.shape[0]:] == -1).all() and (y[b, c.shape[1]:] == -1).all()


def test_lap_rect_is_optimal_under_ties():
    """With exact ties any optimal solution may come back: the extended objective must
    equal lapjv's (sum of matched costs + cost_limit/2 per unmatched row and column)."""
    from oracle import lap
    rng = np.random.default_rng(12)
    costs = _random_costs(rng, 300, -1)
    for limit in (0.4, np.inf):
        x, y = _lap_batch(costs, limit, entry="bm_lap_rect_batch")
        for b, c in enumerate(costs):
            nr, nc = c.shape
            _, xo, _ = lap.lapjv(c, extend_cost=True, cost_limit=limit)
            pad = limit / 2 if limit < np.inf else c.max() + 1

            def obj(xx):
                m = xx >= 0
                return c[np.nonzero(m)[0], xx[m]].sum() + pad * ((nr - m.sum()) + (nc - m.sum()))
            xb = x[b, :nr]
            assert len(set(xb[xb >= 0])) == (xb >= 0).sum()
            for i in range(nr):
                if xb[i] >= 0:
                    assert y[b, xb[i]] == i
            assert abs(obj(xb) - obj(xo)) < 1e-9, "problem %d: %r vs %r" % (b, obj(xb), obj(xo))


def test_lap_batch_empty_problem():
    x, y = _lap_batch([np.zeros((0, 0)), np.ones((2, 2))])
    assert (x[0] == -1).all() and (y[0] == -1).all()


@pytest.mark.parametrize("name", ["iou", "giou", "diou", "ciou", "hmiou", "centroid"])
def test_iou_family_bit_exact(name):
    from boxmot_b200 import _lib
    from oracle import assoc
    rng = np.random.default_rng(3)

    def boxes(n):
        c = rng.random((n, 2)) * [1800, 900]
        wh = rng.uniform(20, 300, (n, 2))
        return np.c_[c, c + wh]

    a, b = boxes(257), boxes(131)
    a[:5] = b[:5]                               # identical boxes -> IoU exactly 1
    d_a, d_b = _dev(a), _dev(b)
    out = torch.zeros((len(a), len(b)), dtype=torch.float64, device="cuda")
    _lib.check(_lib.lib().bm_iou_matrix(_lib.ASSO_FUNCS[name], ctypes.c_void_p(d_a.data_ptr()), len(a),
                                        ctypes.c_void_p(d_b.data_ptr()), len(b), 1920.0, 1080.0,
                                        ctypes.c_void_p(out.data_ptr()), None), "bm_iou_matrix")
    torch.cuda.synchronize()
    want = assoc.asso_func(name, 1920, 1080)(a, b)
    got = out.cpu().numpy()
    if name == "ciou":                           # arctan differs from libm by <= 1 ulp
        np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-14)
    else:
        assert np.array_equal(got, want)


def test_xysr_predict_update_vs_oracle():
    from boxmot_b200 import _lib
    from oracle import kalman
    rng = np.random.default_rng(4)
    n = 300
    P0, Q, R = kalman.xysr_constants(0.01, 0.0001)
    xs = np.zeros((n, 7))
    Ps = np.zeros((n, 7, 7))
    zs = np.zeros((n, 4))
    for i in range(n):
        A = rng.normal(size=(7, 7))
        Ps[i] = A @ A.T + P0 * rng.uniform(0.1, 2)
        xs[i] = np.r_[rng.uniform(100, 1500, 2), rng.uniform(2000, 40000), rng.uniform(0.3, 0.8),
                      rng.normal(0, 3, 3)]
        zs[i] = xs[i, :4] + rng.normal(0, 2, 4) * [1, 1, 50, 0.01]
    d_x, d_P, d_z = _dev(xs), _dev(Ps.reshape(n, 49)), _dev(zs)
    L = _lib.lib()
    _lib.check(L.bm_xysr_predict(ctypes.c_void_p(d_x.data_ptr()), ctypes.c_void_p(d_P.data_ptr()), n,
                                 0.01, 0.0001, None), "predict")
    torch.cuda.synchronize()
    px, pP = d_x.cpu().numpy(), d_P.cpu().numpy().reshape(n, 7, 7)
    for i in range(n):
        wx, wP = kalman.xysr_predict(xs[i], Ps[i], Q)
        assert np.array_equal(px[i], wx) and np.array_equal(pP[i], wP)   # 0/1 matrices: bit exact
    _lib.check(L.bm_xysr_update(ctypes.c_void_p(d_x.data_ptr()), ctypes.c_void_p(d_P.data_ptr()),
                                ctypes.c_void_p(d_z.data_ptr()), n, None), "update")
    torch.cuda.synchronize()
    ux, uP = d_x.cpu().numpy(), d_P.cpu().numpy().reshape(n, 7, 7)
    for i in range(n):
        wx, wP = kalman.xysr_update(px[i], pP[i], zs[i], R)
        np.testing.assert_allclose(ux[i], wx, rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(uP[i], wP, rtol=1e-9, atol=1e-9)


def test_xysr_block_form_is_bit_identical():
    """The engine keeps only the 13 covariance entries that can be non-zero; its predict /
    update must reproduce the dense 7x7 kernels bit for bit over a long random schedule."""
    from boxmot_b200 import _lib
    rng = np.random.default_rng(6)
    n = 512
    z0 = np.c_[rng.uniform(100, 1500, (n, 2)), rng.uniform(2000, 40000, n), rng.uniform(0.3, 0.8, n)]
    x = np.c_[z0, np.zeros((n, 3))]
    Pb = np.tile(np.array([10.0, 0, 0, 1e4] * 3 + [10.0]), (n, 1))
    P = np.zeros((n, 7, 7))
    for b in range(3):
        P[:, b, b], P[:, b + 4, b + 4] = 10.0, 1e4
    P[:, 3, 3] = 10.0
    d_x, d_P = _dev(x), _dev(P.reshape(n, 49))
    d_xb, d_Pb = _dev(x), _dev(Pb)
    L = _lib.lib()
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    for step in range(40):
        _lib.check(L.bm_xysr_predict(p(d_x), p(d_P), n, 0.01, 0.0001, None), "predict")
        _lib.check(L.bm_xysrb_predict(p(d_xb), p(d_Pb), n, 0.01, 0.0001, None), "predict_b")
        if step % 3 != 2:
            z = d_x.cpu().numpy()[:, :4] + rng.normal(0, 2, (n, 4)) * [1, 1, 50, 0.01]
            d_z = _dev(z)
            _lib.check(L.bm_xysr_update(p(d_x), p(d_P), p(d_z), n, None), "update")
            _lib.check(L.bm_xysrb_update(p(d_xb), p(d_Pb), p(d_z), n, None), "update_b")
        torch.cuda.synchronize()
        full, blk = d_P.cpu().numpy().reshape(n, 7, 7), d_Pb.cpu().numpy()
        assert np.array_equal(d_x.cpu().numpy(), d_xb.cpu().numpy()), "step %d x" % step
        exp = np.zeros_like(full)
        for b in range(3):
            exp[:, b, b], exp[:, b, b + 4] = blk[:, 4 * b], blk[:, 4 * b + 1]
            exp[:, b + 4, b], exp[:, b + 4, b + 4] = blk[:, 4 * b + 2], blk[:, 4 * b + 3]
        exp[:, 3, 3] = blk[:, 12]
        assert np.array_equal(full, exp), "step %d P" % step


@pytest.mark.parametrize("kind", ["xyah", "xywh"])
def test_cv8_kalman_kernels_match_reference_vectors(kind):
    """bm_kf_cv8 / bm_kf_cv8_gating (block-form XYAH / XYWH filters) against the reference's own
    KalmanFilterXYAH / XYWH outputs (tests/golden/ref_kf_cv8.npz): initiate, predict, update with
    NSA confidence, multi_predict, gating distances (rows a1-a5).  BIT-IDENTICAL: the block form carries
    both off-diagonal entries of each 2x2 block and reproduces scipy's reciprocal-of-the-diagonal solves."""
    import os
    from boxmot_b200 import _lib
    from conftest import GOLD
    z = np.load(os.path.join(GOLD, "ref_kf_cv8.npz"))
    k = {"xyah": 0, "xywh": 1}[kind]
    means, covs, zs, conf = z[kind + "_means"], z[kind + "_covs"], z[kind + "_zs"], z[kind + "_conf"]
    n = means.shape[1]
    L = _lib.lib()
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    d_mean = torch.zeros((n, 8), dtype=torch.float64, device="cuda")
    d_cov = torch.zeros((n, 8, 8), dtype=torch.float64, device="cuda")
    _lib.check(L.bm_kf_cv8(k, 0, p(d_mean), p(d_cov), p(_dev(z[kind + "_z0"])), None, n, None), "initiate")
    step = 0

    def check():
        torch.cuda.synchronize()
        assert np.array_equal(d_mean.cpu().numpy(), means[step]), step      # bit for bit vs the reference's filter
        assert np.array_equal(d_cov.cpu().numpy(), covs[step]), step

    check()
    for i in range(zs.shape[0]):
        _lib.check(L.bm_kf_cv8(k, 1, p(d_mean), p(d_cov), None, None, n, None), "predict")
        step += 1
        check()
        cf = _dev(conf[i]) if i % 2 else None
        _lib.check(L.bm_kf_cv8(k, 2, p(d_mean), p(d_cov), p(_dev(zs[i])), p(cf), n, None), "update")
        step += 1
        check()
    meas = _dev(z[kind + "_meas"])
    out = torch.zeros((n, meas.shape[0]), dtype=torch.float64, device="cuda")
    for op in (0, 1):
        for gi, metric in enumerate(("maha", "gaussian")):
            _lib.check(L.bm_kf_cv8_gating(k, p(d_mean), p(d_cov), n, p(meas), meas.shape[0], op, gi, p(out), None),
                       "gating")
            torch.cuda.synchronize()
            assert np.array_equal(out.cpu().numpy(), z["%s_gate_%s_%d" % (kind, metric, op)]), (metric, op)
    # multi_predict of the final state (the reference's batched form sums P[i][i]'s four terms in another order)
    _lib.check(L.bm_kf_cv8(k, 3, p(d_mean), p(d_cov), None, None, n, None), "multi_predict")
    torch.cuda.synchronize()
    assert np.array_equal(d_mean.cpu().numpy(), z[kind + "_mp_mean"])
    assert np.array_equal(d_cov.cpu().numpy(), z[kind + "_mp_cov"])


@pytest.mark.parametrize("shape", [(5, 3, 32), (128, 16, 64), (200, 77, 512), (300, 260, 96), (37, 130, 100)])
@pytest.mark.parametrize("mode", [0, 1])
def test_cosine_distance_matches_scipy(shape, mode):
    """bm_cosine_distance (row a14: embedding_distance) against scipy's cdist on the same float32
    features: mode 0 is the tcgen05 / TMEM contraction (3xTF32), mode 1 the fp32 FFMA form.
    Tolerance 2e-6 absolute on a distance in [0, 2] (fp32 accumulation of unit-norm vectors)."""
    from scipy.spatial.distance import cdist
    from boxmot_b200 import _lib
    na, nb, E = shape
    rng = np.random.default_rng(na * 1000 + nb)
    a = rng.normal(0, 1, (na, E)).astype(np.float32)
    b = rng.normal(0, 1, (nb, E)).astype(np.float32)
    b[: min(nb, na)] += 3 * a[: min(nb, na)]          # some close pairs
    a[0] *= 100.0                                      # norms are divided out
    d_a, d_b = _dev(a), _dev(b)
    out = torch.full((na, nb), -7.0, dtype=torch.float32, device="cuda")
    work = torch.zeros((2 * (na + nb),), dtype=torch.int32, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    _lib.check(_lib.lib().bm_cosine_distance(p(d_a), na, p(d_b), nb, E, p(out), p(work), mode, None), "cosine")
    torch.cuda.synchronize()
    want = np.maximum(0.0, cdist(a.astype(np.float64), b.astype(np.float64), "cosine"))
    got = out.cpu().numpy()
    err = np.abs(got - want).max()
    print("cosine mode %d shape %s max abs err %.3g" % (mode, shape, err))
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-6 if mode else 2e-5)


def test_strongsort_nn_cosine_distance_matches_reference_vectors():
    """bm_nn_cosine_distance (tcgen05 contraction per track, minimum over the sample bank in the
    epilogue) against NearestNeighborDistanceMetric.distance of the reference (row a21)."""
    import ctypes, os
    from conftest import GOLD
    from boxmot_b200 import _lib
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))
    bank, nb, feats = z["bank"], z["nb"].astype(np.int32), z["feats"]
    T, B, E = bank.shape
    D = feats.shape[0]

    def run(bank, nb, feats):
        T, B, E = bank.shape
        D = feats.shape[0]
        d_bank, d_nb, d_f = torch.as_tensor(bank).cuda(), torch.as_tensor(nb).cuda(), torch.as_tensor(feats).cuda()
        d_out = torch.full((T, D), -7.0, dtype=torch.float32, device="cuda")
        d_work = torch.zeros((max(B, D) + T * B + D + 16,), dtype=torch.int32, device="cuda")
        rc = _lib.lib().bm_nn_cosine_distance(d_bank.data_ptr(), d_nb.data_ptr(), T, B, d_f.data_ptr(), D, E,
                                              d_out.data_ptr(), d_work.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        return d_out.cpu().numpy().astype(np.float64)

    np.testing.assert_allclose(run(bank, nb, feats), z["nn"], rtol=0, atol=2e-6)
    # larger shapes than the fixture (several row tiles, a bank wider than one column tile, E = 768)
    rng = np.random.default_rng(3)
    T2, B2, E2, D2 = 9, 300, 768, 150
    bank2 = rng.normal(0, 1, (T2, B2, E2)).astype(np.float32)
    nb2 = rng.integers(0, B2 + 1, T2).astype(np.int32)
    nb2[0], nb2[1] = 0, B2
    feats2 = rng.normal(0, 1, (D2, E2)).astype(np.float32)
    want = so.nn_cosine_distance([bank2[t, : nb2[t]].astype(np.float64) for t in range(T2)], feats2.astype(np.float64))
    np.testing.assert_allclose(run(bank2, nb2, feats2), want, rtol=0, atol=3e-6)


def test_strongsort_min_cost_matching_matches_reference_vectors():
    """bm_min_cost_matching_batch against the reference's min_cost_matching (row a18) on the
    fixture's 48 problems, and against the scipy-based oracle on larger random ones."""
    import os
    from conftest import GOLD
    from boxmot_b200 import _lib
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))

    def run(costs, nrs, ncs, maxd):
        P, NR, NC = costs.shape
        d_c = torch.as_tensor(np.ascontiguousarray(costs)).cuda()
        d_nr, d_nc = torch.as_tensor(nrs.astype(np.int32)).cuda(), torch.as_tensor(ncs.astype(np.int32)).cuda()
        d_x = torch.zeros((P, NR), dtype=torch.int32, device="cuda")
        d_y = torch.zeros((P, NC), dtype=torch.int32, device="cuda")
        rc = _lib.lib().bm_min_cost_matching_batch(d_c.data_ptr(), P, NR, NC, d_nr.data_ptr(), d_nc.data_ptr(),
                                                   float(maxd), d_x.data_ptr(), d_y.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        return d_x.cpu().numpy(), d_y.cpu().numpy()

    x, y = run(z["costs"], z["nrs"], z["ncs"], float(z["max_distance"]))
    assert np.array_equal(x, z["xs"])
    for p in range(len(x)):
        for r, c in enumerate(x[p]):
            if c >= 0:
                assert y[p, c] == r
    rng = np.random.default_rng(8)
    P, NR, NC = 64, 120, 90
    costs = rng.random((P, NR, NC))
    nrs, ncs = rng.integers(0, NR + 1, P), rng.integers(0, NC + 1, P)
    x, _ = run(costs, nrs, ncs, 0.2)
    for p in range(P):
        m, _, _ = so.min_cost_matching(costs[p, : nrs[p], : ncs[p]], 0.2)
        want = -np.ones(NR, dtype=np.int64)
        for r, c in m:
            want[r] = c
        assert np.array_equal(x[p], want), p


def test_strongsort_cost_postprocessing_matches_reference_vectors():
    """bm_strongsort_gated_metric (row a22: gate_cost_matrix + id-preservation weighting),
    bm_strongsort_enhance_cost and bm_strongsort_iou_cost against the reference's own functions run
    on its own Track / Detection objects (fixture), and against the oracle on a larger random
    problem.  fp64 throughout; 1e-12 relative (the gate: sqrt/division order of the Cholesky solve)."""
    import os
    from conftest import GOLD
    from boxmot_b200 import _lib
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))

    def dev(a):
        return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).cuda()

    def run(cost0, means, covs, trk, tlwh, det, lam, idw):
        T, D = cost0.shape
        xyah = tlwh.copy()
        xyah[:, :2] += xyah[:, 2:] / 2
        xyah[:, 2] /= xyah[:, 3]
        d_m, d_c, d_t, d_z, d_l, d_d = dev(means), dev(covs), dev(trk), dev(xyah), dev(tlwh), dev(det)
        g, e = dev(cost0), dev(cost0)
        i = torch.zeros((T, D), dtype=torch.float64, device="cuda")
        L = _lib.lib()
        assert L.bm_strongsort_gated_metric(g.data_ptr(), d_m.data_ptr(), d_c.data_ptr(), d_t.data_ptr(), T,
                                            d_z.data_ptr(), D, float(lam), float(idw), None) == 0
        assert L.bm_strongsort_enhance_cost(e.data_ptr(), d_t.data_ptr(), T, d_d.data_ptr(), D, None) == 0
        assert L.bm_strongsort_iou_cost(d_m.data_ptr(), d_t.data_ptr(), T, d_l.data_ptr(), D, i.data_ptr(), None) == 0
        torch.cuda.synchronize()
        return g.cpu().numpy(), e.cpu().numpy(), i.cpu().numpy(), xyah

    det = np.c_[z["g_dq"], z["g_dconf"], z["g_dcls"]]
    g, e, i, _ = run(z["g_cost0"], z["g_means"], z["g_covs"], z["g_trk"], z["g_tlwh"], det, z["g_lambda"], z["g_idw"])
    np.testing.assert_allclose(g, z["g_gated"], rtol=1e-12, atol=0)
    assert np.array_equal(g > 500, z["g_gated"] > 500) and (g < 1.0).sum() > 20
    np.testing.assert_allclose(e, z["g_enh"], rtol=1e-14, atol=0)
    np.testing.assert_allclose(i, z["g_iou"], rtol=1e-13, atol=1e-15)

    # larger random problem against the oracle: 300 tracks x 400 detections
    rng = np.random.default_rng(17)
    T, D = 300, 400
    from oracle import kalman as K
    means, covs = np.zeros((T, 8)), np.zeros((T, 8, 8))
    for t in range(T):
        wh = rng.uniform(20, 200, 2)
        m, c = K.cv8_initiate("xyah", np.r_[rng.random(2) * [1800, 1000], wh[0] / wh[1], wh[1]])
        for _ in range(int(rng.integers(1, 6))):
            m, c = K.cv8_predict("xyah", m, c)
        means[t], covs[t] = m, c
    trk = np.zeros((T, so.TRK_FIELDS))
    trk[:, so.TRK_AGE] = rng.integers(1, 40, T)
    trk[:, so.TRK_HITS] = rng.integers(1, 30, T)
    trk[:, so.TRK_TSU] = rng.integers(0, 4, T)
    trk[:, so.TRK_MOTION] = rng.random(T)
    trk[:, so.TRK_APPEAR] = rng.random(T)
    trk[:, so.TRK_QUALITY] = rng.random(T)
    trk[:, so.TRK_HI_STREAK] = rng.integers(0, 7, T)
    trk[:, so.TRK_LO_STREAK] = rng.integers(0, 5, T)
    trk[:, so.TRK_HAS_VEL] = rng.integers(0, 2, T)
    trk[:, so.TRK_VX:so.TRK_VY + 1] = rng.normal(0, 3, (T, 2))
    trk[:, so.TRK_CONF] = rng.random(T)
    trk[:, so.TRK_CLS] = rng.integers(0, 3, T)
    tlwh = np.zeros((D, 4))
    for d in range(D):
        m = means[d % T]
        h = m[3] * (1 + rng.normal(0, 0.04))
        w = m[2] * m[3] * (1 + rng.normal(0, 0.04))
        c = m[:2] + rng.normal(0, 0.05, 2) * m[3]
        tlwh[d] = [c[0] - w / 2, c[1] - h / 2, w, h]
    det = np.c_[rng.random(D) * 1.2, rng.random(D), rng.integers(0, 3, D)]
    cost0 = rng.uniform(0, 0.5, (T, D))
    g, e, i, xyah = run(cost0, means, covs, trk, tlwh, det, 0.995, 0.1)
    want = so.gated_metric(cost0, means, covs, trk, xyah, 0.995, 0.1)
    assert np.array_equal(g > 500, want > 500) and 100 < (want < 500).sum()
    np.testing.assert_allclose(g, want, rtol=1e-12, atol=0)
    np.testing.assert_allclose(e, so.enhance_cost_matrix(cost0, trk, det[:, 0], det[:, 1], det[:, 2]), rtol=1e-14)
    np.testing.assert_allclose(i, so.iou_cost(means, trk[:, so.TRK_TSU], tlwh), rtol=1e-13, atol=1e-15)


def test_strongsort_feature_update_matches_reference_vectors():
    """bm_strongsort_feature_update (row a16, StrongSORT form) against the features the reference's
    Track.update appended (fixture), incl. tracks without a stored feature, and against the oracle on
    768-d CLIPReID-shaped vectors (BASELINE configs[3])."""
    import os
    from conftest import GOLD
    from boxmot_b200 import _lib
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))

    def run(last, has, feat, conf, ac0, alpha):
        n, E = feat.shape
        d_l = torch.as_tensor(np.ascontiguousarray(last, dtype=np.float32)).cuda()
        d_h = torch.as_tensor(np.ascontiguousarray(has, dtype=np.int32)).cuda()
        d_f = torch.as_tensor(np.ascontiguousarray(feat, dtype=np.float32)).cuda()
        d_c = torch.as_tensor(np.ascontiguousarray(conf, dtype=np.float64)).cuda()
        d_a = torch.as_tensor(np.ascontiguousarray(ac0, dtype=np.float64)).cuda()
        d_s = torch.zeros((n,), dtype=torch.float64, device="cuda")
        d_o = torch.zeros((n, E), dtype=torch.float32, device="cuda")
        assert _lib.lib().bm_strongsort_feature_update(d_l.data_ptr(), d_h.data_ptr(), d_f.data_ptr(), d_c.data_ptr(), n, E,
                                                       float(alpha), d_a.data_ptr(), d_s.data_ptr(), d_o.data_ptr(),
                                                       None) == 0
        torch.cuda.synchronize()
        return d_o.cpu().numpy(), d_a.cpu().numpy(), d_s.cpu().numpy()

    n = len(z["f_out"])
    out, ac, sim = run(z["f_last"], np.ones(n), z["f_feat"], z["f_conf"], z["f_ac0"], z["f_alpha"])
    np.testing.assert_allclose(out, z["f_out"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(ac, z["f_ac1"], rtol=0, atol=1e-6)
    rng = np.random.default_rng(23)
    N, E = 513, 768
    ident = rng.normal(0, 1, (N, E))
    last = ident / np.linalg.norm(ident, axis=1, keepdims=True)
    feat = (ident + rng.choice([0.05, 0.8, 3.0], (N, 1)) * rng.normal(0, 1, (N, E))) * rng.uniform(0.5, 4, (N, 1))
    conf, ac0 = rng.uniform(0.05, 0.99, N), rng.uniform(0.3, 1.0, N)
    has = (np.arange(N) % 7 != 0).astype(np.int32)
    last32, feat32 = last.astype(np.float32), feat.astype(np.float32)
    out, ac, sim = run(last32, has, feat32, conf, ac0, 0.9)
    for i in range(N):
        w_out, w_ac, w_sim = so.feature_update(last32[i] if has[i] else None, feat32[i], conf[i], ac0[i], 0.9)
        if has[i] and min(abs(w_sim - 0.7), abs(w_sim - 0.4)) < 1e-6:
            continue                      # branch decided by rounding
        np.testing.assert_allclose(out[i], w_out, rtol=0, atol=2e-7)
        assert abs(ac[i] - w_ac) < 1e-12 and (np.isnan(sim[i]) if not has[i] else abs(sim[i] - w_sim) < 1e-12)


def test_strongsort_track_state_machine_matches_reference_vectors():
    """bm_strongsort_track (Track.predict / update / mark_missed / camera_update, track.py:173-317) against every call
    recorded from the reference's own Track objects: Kalman state 1e-9 relative (block-form filter vs
    scipy's Cholesky solve), table columns 1e-12, the state column exact."""
    import os
    from conftest import GOLD
    from boxmot_b200 import _lib
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))
    ops, par = z["t_op"], z["t_par"].astype(int)
    checked = 0
    for op in (0, 1, 2, 3):
        keys = sorted(set((int(p[0]), int(p[1])) for p in par[ops == op]))
        for n_init, max_age in keys:
            sel = np.where((ops == op) & (par[:, 0] == n_init) & (par[:, 1] == max_age))[0]
            d_m = torch.as_tensor(np.ascontiguousarray(z["t_m0"][sel])).cuda()
            d_c = torch.as_tensor(np.ascontiguousarray(z["t_c0"][sel])).cuda()
            d_a = torch.as_tensor(np.ascontiguousarray(z["t_a0"][sel])).cuda()
            d_d = torch.as_tensor(np.ascontiguousarray(z["t_det"][sel])).cuda()
            assert d_a.shape[1] == so.TS_FIELDS == 42
            assert _lib.lib().bm_strongsort_track(op, d_m.data_ptr(), d_c.data_ptr(), d_a.data_ptr(), d_d.data_ptr(),
                                                  len(sel), n_init, max_age, None) == 0
            torch.cuda.synchronize()
            a = d_a.cpu().numpy()
            assert np.array_equal(a[:, so.TS_STATE], z["t_a1"][sel][:, so.TS_STATE])
            np.testing.assert_allclose(a, z["t_a1"][sel], rtol=1e-12, atol=1e-12)
            np.testing.assert_allclose(d_m.cpu().numpy(), z["t_m1"][sel], rtol=1e-9, atol=1e-9)
            np.testing.assert_allclose(d_c.cpu().numpy(), z["t_c1"][sel], rtol=1e-9, atol=1e-9)
            checked += len(sel)
    assert checked == len(ops) > 4000


def test_strongsort_matching_cascade_matches_reference_vectors():
    """boxmot_b200.strongsort_units.matching_cascade (the cascade composed of the C-ABI units: tcgen05
    sample-bank distance -> gate / blend / id preservation -> cost enhancement -> thresholded exact
    assignment, level by level) against the reference's matching_cascade on its own objects: the same
    matches, unmatched tracks and unmatched detections."""
    import os
    from conftest import GOLD
    from boxmot_b200.strongsort_units import matching_cascade
    with np.load(os.path.join(GOLD, "ref_strongsort_units.npz")) as f:
        z = {k: f[k] for k in f.files if k.startswith("c_")}
    tlwh = z["c_tlwh"]
    xyah = tlwh.copy()
    xyah[:, :2] += xyah[:, 2:] / 2
    xyah[:, 2] /= xyah[:, 3]
    m, ut, ud = matching_cascade(z["c_means"], z["c_covs"], z["c_trk"], z["c_bank"], z["c_nb"], xyah, z["c_dattr"],
                                 z["c_feats"], range(len(z["c_nb"])), range(len(tlwh)), float(z["c_maxd"]),
                                 int(z["c_depth"]))
    assert sorted(m) == [tuple(p) for p in z["c_matches"]] and len(m) >= 15
    assert sorted(ut) == list(z["c_ut"]) and sorted(ud) == list(z["c_ud"])
