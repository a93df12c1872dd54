"""CPU suite: the oracle against the reference-generated golden fixtures, the lapjv
restatement against scipy, and the C-ABI library's exports (no compute without a GPU)."""
import ctypes
import os

import numpy as np
import pytest
from scipy.optimize import linear_sum_assignment

from conftest import (BOOSTTRACK_CASES, BOOSTTRACK_YAML, boosttrack_case, DEEPOCSORT_CASES, DEEPOCSORT_YAML, deepocsort_case, BOTSORT_CASES, BOTSORT_YAML, BYTETRACK_YAML, OCSORT_YAML, ROOT, botsort_case, assert_rows_match, case_frames, golden_rows,
                      split_by_frame)
from oracle import lap
from oracle.ocsort import OcSortOracle


def test_lapjv_optimum_matches_scipy():
    rng = np.random.default_rng(0)
    for it in range(200):
        nr, nc = rng.integers(1, 40, 2)
        c = rng.random((nr, nc))
        opt, x, y = lap.lapjv(c, extend_cost=True)
        r, cc = linear_sum_assignment(c)
        assert abs(opt - c[r, cc].sum()) < 1e-9
        assert (x >= 0).sum() == min(nr, nc)
        L = 0.5
        _, x, y = lap.lapjv(c, extend_cost=True, cost_limit=L)
        n = nr + nc
        e = np.full((n, n), L / 2)
        e[nr:, nc:] = 0
        e[:nr, :nc] = c
        r, cc = linear_sum_assignment(e)
        mine = sum(c[i, x[i]] for i in range(nr) if x[i] >= 0) + L / 2 * ((x < 0).sum() + (y < 0).sum())
        assert abs(e[r, cc].sum() - mine) < 1e-9
        for i in range(nr):
            if x[i] >= 0:
                assert y[x[i]] == i and c[i, x[i]] <= L


def test_lapjv_skip_rule_is_result_neutral():
    """oracle_lapjv_skip models the CUDA solver's one shortcut; it must equal the plain
    algorithm on tie-heavy inputs."""
    lib = lap._lib()
    lib.oracle_lapjv_skip.restype = ctypes.c_int
    lib.oracle_lapjv_skip.argtypes = lib.oracle_lapjv.argtypes
    rng = np.random.default_rng(1)
    for it in range(600):
        nr, nc = rng.integers(1, 70, 2)
        c = rng.random((nr, nc))
        if it % 3 == 1:
            c = np.round(c * 4) / 4
        if it % 3 == 2:
            c = -np.where(rng.random((nr, nc)) < 0.9, 0.0, c)
        lim = 1.7976931348623157e308 if it % 2 else 0.4
        c = np.ascontiguousarray(c)
        res = []
        for fn in (lib.oracle_lapjv, lib.oracle_lapjv_skip):
            x = np.empty(nr, np.int32)
            y = np.empty(nc, np.int32)
            fn(c.ctypes.data, nr, nc, 1, lim, x.ctypes.data, y.ctypes.data, None)
            res.append((x, y))
        assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])


@pytest.mark.parametrize("case", ["synth_k100", "synth_k30_byte", "synth_k30_giou",
                                  "synth_k50_dt1", "synth_k40_ciou", "MOT17_02_FRCNN"])
def test_ocsort_oracle_matches_reference_golden(case):
    want, kw = golden_rows("ocsort", case)
    frames = case_frames(case)
    ora = OcSortOracle(**{**OCSORT_YAML, **kw})
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = ora.update(d)
        assert_rows_match(got, want_f[f], f, box_rtol=1e-12)


@pytest.mark.parametrize("case", ["synth_k100", "synth_k30_byte", "synth_k30_giou", "synth_k50_dt1",
                                  "synth_k40_ciou", "MOT17_02_FRCNN", "MOT17_04_FRCNN"])
def test_bytetrack_oracle_matches_reference_golden(case):
    """BASELINE configs[0] (MOT17-mini, bytetrack.yaml) and the synthetic streams."""
    from oracle.bytetrack import ByteTrackOracle
    want, _ = golden_rows("bytetrack", case)
    frames = case_frames(case)
    ora = ByteTrackOracle(**BYTETRACK_YAML)
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = ora.update(d)
        assert_rows_match(got, want_f[f], f, box_rtol=1e-12)


@pytest.mark.parametrize("case", [c for c in BOTSORT_CASES if c != "MOT17_04_FRCNN"])   # MOT17-04: GPU tests (CPU suite time)
def test_botsort_oracle_matches_reference_golden(case):
    """BoT-SORT (BASELINE configs[2] tracker): XYWH filter, IoU / embedding fusion, feature EMA."""
    from oracle.botsort import BotSortOracle
    from conftest import botsort_warps
    frames, embs, kw, want = botsort_case(case)
    warps = botsort_warps(case)
    ora = BotSortOracle(**{**BOTSORT_YAML, **kw})
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = ora.update(d, None if embs is None else embs[f].copy(), None if warps is None else warps[f])
        assert_rows_match(got, want_f[f], f, box_rtol=1e-9)


@pytest.mark.parametrize("case", [c for c in DEEPOCSORT_CASES if c != "MOT17_04_FRCNN"])   # MOT17-04: GPU tests (CPU suite time)
def test_deepocsort_oracle_matches_reference_golden(case):
    """DeepOCSORT: OC-SORT + embedding term with the adaptive top-2 weighting + CMC correction
    (incl. the reference's doubly-warped newest observation)."""
    from oracle.deepocsort import DeepOcSortOracle
    frames, embs, warps, kw, want = deepocsort_case(case)
    ora = DeepOcSortOracle(**{**DEEPOCSORT_YAML, **kw})
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = ora.update(d, None, embs[f].copy(), warp=None if warps is None else warps[f])
        assert_rows_match(got, want_f[f], f, box_rtol=1e-12)


def test_deepocsort_oracle_per_class_matches_reference_golden():
    """per_class=True: one oracle per class id, ids renumbered from the shared creation counter
    (basetracker.py:155-201, deepocsort.py:56-58) -- the same host-side logic as the adapter."""
    from conftest import deepocsort_per_class_case
    from oracle.deepocsort import DeepOcSortOracle
    frames, embs, warps, want = deepocsort_per_class_case()
    oras = [DeepOcSortOracle(**DEEPOCSORT_YAML) for _ in range(3)]
    gids, count = [[] for _ in oras], 1
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        rows = []
        for c, ora in enumerate(oras):
            sel = d[:, 5] == c
            r = ora.update(d[sel], None, embs[f][sel].copy(), warp=warps[f])
            while len(gids[c]) < ora.next_id - 1:
                gids[c].append(count)
                count += 1
            if len(r):
                r = r.copy()
                r[:, 4] = [gids[c][int(i) - 1] for i in r[:, 4]]
                rows.append(r)
        got = np.vstack(rows) if rows else np.empty((0, 8))
        assert_rows_match(got, want_f[f], f, box_rtol=1e-12)


@pytest.mark.parametrize("case", [c for c in BOOSTTRACK_CASES if c != "MOT17_04_FRCNN"])   # MOT17-04: GPU tests (CPU suite time)
def test_boosttrack_oracle_matches_reference_golden(case):
    """BoostTrack(++): 8-state filter, DLO / DUO confidence boosts, soft-BIoU / Mahalanobis / shape /
    embedding cost, camera_update with similarity warps."""
    from oracle.boosttrack import BoostTrackOracle
    frames, embs, warps, kw, want = boosttrack_case(case)
    ora = BoostTrackOracle(**{**BOOSTTRACK_YAML, **kw})
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = ora.update(d, None, embs[f].copy(), warp=None if warps is None else warps[f])
        assert_rows_match(got, want_f[f], f, box_rtol=1e-9, conf_rtol=1e-12)


def test_strongsort_unit_oracle_matches_reference_vectors():
    """oracle/strongsort.py against vectors produced by the reference's own
    NearestNeighborDistanceMetric.distance and min_cost_matching."""
    from conftest import GOLD
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))
    banks = [z["bank"][t, : z["nb"][t]] for t in range(len(z["nb"]))]
    np.testing.assert_allclose(so.nn_cosine_distance(banks, z["feats"]), z["nn"], rtol=0, atol=2e-6)
    for p in range(len(z["nrs"])):
        nr, nc = int(z["nrs"][p]), int(z["ncs"][p])
        m, um_r, um_c = so.min_cost_matching(z["costs"][p, :nr, :nc], float(z["max_distance"]))
        x = -np.ones(z["xs"].shape[1], dtype=np.int64)
        for r, c in m:
            x[r] = c
        assert np.array_equal(x, z["xs"][p])
        assert sorted(um_r) == [r for r in range(nr) if x[r] < 0]


def test_strongsort_gate_oracle_matches_reference_vectors():
    """gated_metric's post-processing (gate_cost_matrix + id-preservation weighting, row a22),
    _enhance_cost_matrix and iou_cost against the reference's own functions run on the reference's
    own Track / Detection objects (oracle/gen_golden.py strongsort_gate_vectors).  Tolerance 1e-12
    relative: pure fp64 arithmetic in the same order."""
    from conftest import GOLD
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))
    tlwh = z["g_tlwh"]
    xyah = tlwh.copy()
    xyah[:, :2] += xyah[:, 2:] / 2
    xyah[:, 2] /= xyah[:, 3]
    g = so.gated_metric(z["g_cost0"], z["g_means"], z["g_covs"], z["g_trk"], xyah, float(z["g_lambda"]),
                        float(z["g_idw"]))
    np.testing.assert_allclose(g, z["g_gated"], rtol=1e-12, atol=0)
    assert (g < 1.0).sum() > 20 and (g > 500).sum() > 1000          # both sides of the gate are exercised
    e = so.enhance_cost_matrix(z["g_cost0"], z["g_trk"], z["g_dq"], z["g_dconf"], z["g_dcls"])
    np.testing.assert_allclose(e, z["g_enh"], rtol=1e-14, atol=0)
    i = so.iou_cost(z["g_means"], z["g_trk"][:, so.TRK_TSU], tlwh)
    np.testing.assert_allclose(i, z["g_iou"], rtol=1e-13, atol=1e-15)
    assert (i == so.INFTY_COST).any() and (i < 1).sum() > 50


def test_strongsort_feature_update_oracle_matches_reference_vectors():
    """The feature EMA of the reference's Track.update (row a16, StrongSORT form) on vectors harvested
    from reference Track objects: every conf / similarity branch of the adaptive alpha.  1e-6: the
    reference itself mixes float32 and float64 here."""
    from conftest import GOLD
    from oracle import strongsort as so
    z = np.load(os.path.join(GOLD, "ref_strongsort_units.npz"))
    sims = []
    for i in range(len(z["f_out"])):
        sm, ac, sim = so.feature_update(z["f_last"][i], z["f_feat"][i], float(z["f_conf"][i]), float(z["f_ac0"][i]),
                                        float(z["f_alpha"]))
        np.testing.assert_allclose(sm, z["f_out"][i], rtol=0, atol=1e-6)
        assert abs(ac - z["f_ac1"][i]) < 1e-6
        sims.append(sim)
    sims = np.array(sims)
    assert (sims > 0.7).sum() > 30 and ((sims > 0.4) & (sims <= 0.7)).sum() > 30 and (sims <= 0.4).sum() > 30
    assert np.minimum(np.abs(sims - 0.7), np.abs(sims - 0.4)).min() > 1e-5      # no branch decided by rounding


def _cascade_inputs(z):
    tlwh = z["c_tlwh"]
    xyah = tlwh.copy()
    xyah[:, :2] += xyah[:, 2:] / 2
    xyah[:, 2] /= xyah[:, 3]
    return xyah


def test_strongsort_matching_cascade_oracle_matches_reference_vectors():
    """matching_cascade with the gated appearance metric (rows a18 / a21 / a22 chained) against the
    reference's own function on its own Track / Detection objects and sample bank."""
    from conftest import GOLD
    from oracle import strongsort as so
    with np.load(os.path.join(GOLD, "ref_strongsort_units.npz")) as f:
        z = {k: f[k] for k in f.files if k.startswith("c_")}
    banks = [z["c_bank"][t, : z["c_nb"][t]] for t in range(len(z["c_nb"]))]
    m, ut, ud = so.matching_cascade(z["c_means"], z["c_covs"], z["c_trk"], banks, _cascade_inputs(z), z["c_dattr"],
                                    z["c_feats"], range(len(banks)), range(len(z["c_tlwh"])), float(z["c_maxd"]),
                                    int(z["c_depth"]))
    assert sorted(m) == [tuple(p) for p in z["c_matches"]] and len(m) >= 15
    assert sorted(ut) == list(z["c_ut"]) and sorted(ud) == list(z["c_ud"])


def test_strongsort_track_oracle_matches_reference_vectors():
    """Track.predict / update / mark_missed (track.py:195-317) restated on a flat table against every
    call recorded from 64 reference Track objects (4575 calls incl. camera_update): Kalman state, histories, scores,
    streaks and the state transitions.  1e-12."""
    from conftest import GOLD
    from oracle import strongsort as so
    with np.load(os.path.join(GOLD, "ref_strongsort_units.npz")) as f:
        z = {k: f[k] for k in f.files if k.startswith("t_")}          # decompress once, not per access
    ops = z["t_op"]
    assert np.bincount(ops).min() > 300 and set(z["t_a1"][:, so.TS_STATE]) == {1.0, 2.0, 3.0}
    for i in range(len(ops)):
        m, c, a = z["t_m0"][i].copy(), z["t_c0"][i].copy(), z["t_a0"][i].copy()
        n_init, max_age = (int(v) for v in z["t_par"][i])
        if ops[i] == 0:
            m, c = so.track_predict(m, c, a)
        elif ops[i] == 1:
            d = z["t_det"][i]
            m, c = so.track_update(m, c, a, d[:4], d[4], d[5], d[6], n_init)
        elif ops[i] == 2:
            so.track_mark_missed(a, max_age)
        else:
            m = so.track_camera_update(m, a, z["t_det"][i][:6])
        np.testing.assert_allclose(m, z["t_m1"][i], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(c, z["t_c1"][i], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(a, z["t_a1"][i], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("kind", ["xyah", "xywh"])
def test_cv8_kalman_oracle_matches_reference_vectors(kind):
    """oracle/kalman.py cv8_* against known-answer vectors generated from the reference's
    KalmanFilterXYAH / KalmanFilterXYWH (oracle/gen_golden.py --kf): SURVEY rows a1-a5."""
    from conftest import GOLD
    from oracle import kalman as K
    z = np.load(os.path.join(GOLD, "ref_kf_cv8.npz"))
    means, covs, zs, conf = z[kind + "_means"], z[kind + "_covs"], z[kind + "_zs"], z[kind + "_conf"]
    n = means.shape[1]
    cur = [K.cv8_initiate(kind, z0) for z0 in z[kind + "_z0"]]
    step = 0

    def check():
        np.testing.assert_allclose(np.array([m for m, _ in cur]), means[step], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(np.array([c for _, c in cur]), covs[step], rtol=1e-11, atol=1e-13)

    check()
    for k in range(zs.shape[0]):
        cur = [K.cv8_predict(kind, m, c) for m, c in cur]
        step += 1
        check()
        cur = [K.cv8_update(kind, m, c, zs[k, i], conf[k, i] if k % 2 else 0.0) for i, (m, c) in enumerate(cur)]
        step += 1
        check()
    mm, cc = K.cv8_multi_predict(kind, means[-1], covs[-1])
    np.testing.assert_allclose(mm, z[kind + "_mp_mean"], rtol=1e-12)
    np.testing.assert_allclose(cc, z[kind + "_mp_cov"], rtol=1e-11, atol=1e-13)
    for op in (0, 1):
        for metric in ("maha", "gaussian"):
            got = np.array([K.cv8_gating_distance(kind, means[-1][i], covs[-1][i], z[kind + "_meas"], bool(op), metric)
                            for i in range(n)])
            np.testing.assert_allclose(got, z["%s_gate_%s_%d" % (kind, metric, op)], rtol=1e-10)


def test_library_exports_every_declared_symbol():
    from boxmot_b200 import _lib, build
    path = build.build()
    L = ctypes.CDLL(path)
    header = open(os.path.join(ROOT, "include", "boxmot_b200.h")).read()
    import re
    declared = set(re.findall(r"\b(bm_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(L, name), "missing export " + name
    assert set(_lib.EXPORTS) == declared


@pytest.mark.parametrize("case", ["synth_k20_e32", "synth_k25_e96_age8", "synth_k15_e32_ninit3"])
def test_strongsort_oracle_matches_reference_golden(case):
    """The end-to-end StrongSORT oracle (oracle/strongsort_tracker.py) against rows produced by the
    imported reference (R4 / M1 / M2 / R5): sample lists far above nn_budget, id recovery, tentative
    survivors, empty frames.  The remaining four fixtures are checked through the GPU tests (the
    oracle matches all seven: oracle/gen_golden.py --strongsort, tools/check_strongsort.py)."""
    from conftest import strongsort_case
    from oracle.strongsort_tracker import StrongSortOracle
    frames, embs, warps, kw, want = strongsort_case(case)
    o = StrongSortOracle(**kw)
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = o.update(d, embs[f], None if warps is None else warps[f])
        assert_rows_match(got, want_f[f], f, box_rtol=1e-9)


def test_ddot_restatement_equals_numpy_on_the_authoring_cpu():
    """oracle/numerics.c restates OpenBLAS' SkylakeX ddot order (what decides partial_fit's ranking of
    unit vectors).  Where numpy dispatches that kernel -- the authoring container -- np.dot and the
    restatement must agree to the last bit; elsewhere the reference itself would rank differently."""
    threadpoolctl = pytest.importorskip("threadpoolctl")
    arch = {i.get("architecture") for i in threadpoolctl.threadpool_info() if i.get("user_api") == "blas"}
    if arch != {"SkylakeX"}:
        pytest.skip("numpy's BLAS dispatches %s, not the SkylakeX kernels" % sorted(arch))
    rng = np.random.default_rng(0)
    for n in (32, 64, 96, 128, 512, 768, 100, 37, 1000):
        for _ in range(300):
            x = rng.normal(0, 1, n)
            x /= np.linalg.norm(x)
            y = rng.normal(0, 1, n)
            assert lap.ddot(x, x) == float(np.dot(x, x)) and lap.ddot(x, y) == float(np.dot(x, y))
            assert lap.norm(x) == float(np.linalg.norm(x))


def test_iou_family_oracle_equals_reference_functions():
    """oracle/assoc.py's six association functions against boxmot.utils.iou.AssociationFunction of the
    imported reference (authoring container only: the GPU box has no /root/reference)."""
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("reference tree not present")
    ref_shim.load()
    from boxmot.utils.iou import AssociationFunction
    from oracle import assoc
    rng = np.random.default_rng(5)
    a = rng.uniform(0, 900, (40, 2))
    a = np.c_[a, a + rng.uniform(5, 300, (40, 2))]
    b = rng.uniform(0, 900, (55, 2))
    b = np.c_[b, b + rng.uniform(5, 300, (55, 2))]
    b[:10] = a[:10] + rng.normal(0, 3, (10, 4))
    b[10] = a[10]
    for name in ("iou", "giou", "diou", "ciou", "hmiou", "centroid"):
        ref = AssociationFunction(w=1920, h=1080, asso_mode=name).asso_func(a, b)
        mine = assoc.asso_func(name, 1920, 1080)(a, b)
        assert np.array_equal(np.asarray(ref), np.asarray(mine)), name


# ---- per_class for the trackers whose classes the decorator couples (row a25)
def _per_class_frames(with_embs):
    from boxmot_b200.synth import make_stream
    if with_embs:
        from conftest import deepocsort_per_class_case
        fr, em, _, _ = deepocsort_per_class_case()
    else:
        fr, em = make_stream(5, 40, 60), None
    out = []
    for d in fr:
        d = d.copy()
        d[:, 5] = np.floor((d[:, 0] + d[:, 2]) / 2.0 / 640.0).clip(0, 2)
        out.append(d.astype(np.float32).astype(np.float64))
    return out, em


@pytest.mark.parametrize("tracker,case", [("bytetrack", "synth_k40_perclass"), ("botsort", "synth_k40_e32_perclass"),
                                          ("boosttrack", "synth_k40_perclass"),
                                          ("boosttrack", "synth_k40_perclass_age200")])
def test_per_class_oracles_match_reference_golden(tracker, case):
    """oracle/per_class.py (the decorator's loop over the numpy oracles: active list swapped, lost list / BoostTrack
    pool shared, frame counter rewound) against the reference's own per_class=True output."""
    from conftest import BOOSTTRACK_YAML, BOTSORT_YAML, BYTETRACK_YAML, assert_rows_match, golden_rows, split_by_frame
    from oracle.per_class import per_class_update
    want, kw = golden_rows(tracker, case)
    kw = {k: v for k, v in kw.items() if k != "per_class"}
    frames, embs = _per_class_frames(tracker != "bytetrack")
    if tracker == "bytetrack":
        from oracle.bytetrack import ByteTrackOracle
        o = ByteTrackOracle(**BYTETRACK_YAML)
    elif tracker == "botsort":
        from oracle.botsort import BotSortOracle
        o = BotSortOracle(**BOTSORT_YAML)
    else:
        from oracle.boosttrack import BoostTrackOracle
        o = BoostTrackOracle(**dict(BOOSTTRACK_YAML, **kw))
    wf = split_by_frame(want, len(frames))
    n = 0
    for f, d in enumerate(frames):
        got = per_class_update(o, d, None if embs is None else embs[f].astype(np.float64), swap=tracker != "boosttrack")
        assert_rows_match(got, wf[f], f, conf_rtol=1e-9 if tracker == "boosttrack" else 0.0)
        n += len(got)
    assert n == len(want)
