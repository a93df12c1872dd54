"""End-to-end GPU parity of the DeepOCSORT engine (OC-SORT + embedding term with the adaptive
top-2 weighting on the tensor-core contraction + embedding EMA + CMC correction) against
reference-generated golden fixtures and the oracle, through the reference-facing API and the
batched engine (C-ABI bm_deepocsort_*)."""
import numpy as np
import pytest

from conftest import (DEEPOCSORT_CASES, DEEPOCSORT_YAML, assert_rows_match, deepocsort_case, deepocsort_per_class_case,
                      split_by_frame)

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
IMG = np.zeros((1080, 1920, 3), np.uint8)


@pytest.mark.parametrize("case", list(DEEPOCSORT_CASES))
def test_deepocsort_tracker_matches_reference_golden(case):
    """create_tracker("deepocsort").update(dets, img, embs[, warp]) row-for-row against the
    reference: ids / conf / cls / det_ind and row order exact, boxes within 1e-6."""
    import boxmot_b200
    frames, embs, warps, kw, want = deepocsort_case(case)
    trk = boxmot_b200.create_tracker("deepocsort", boxmot_b200.get_tracker_config("deepocsort"),
                                     evolve_param_dict={**DEEPOCSORT_YAML, **kw})
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = trk.update(d, IMG, embs[f].copy(), warp=None if warps is None else warps[f])
        assert_rows_match(got, want_f[f], f)


def test_deepocsort_per_class_matches_reference_golden():
    """per_class=True (basetracker.py:155-201): one track list per class id, the frame's warp applied
    to each, det_ind relative to the class subset, ONE id counter walked in class order."""
    import boxmot_b200
    frames, embs, warps, want = deepocsort_per_class_case()
    trk = boxmot_b200.create_tracker("deepocsort", boxmot_b200.get_tracker_config("deepocsort"), per_class=True)
    want_f = split_by_frame(want, len(frames))
    seen = 0
    for f, d in enumerate(frames):
        got = trk.update(d, IMG, embs[f].copy(), warp=warps[f])
        assert_rows_match(got, want_f[f], f)
        seen += len(got)
    assert seen == len(want) and len(np.unique(want[:, 7])) == 3


@pytest.mark.parametrize("aw_off", [False, True])
def test_deepocsort_engine_batched_streams_match_oracle(aw_off):
    """32 streams with 64-d embeddings and per-stream scale/shift warps stepped together; every
    stream equals a private oracle, including the Kalman state."""
    from boxmot_b200.engine import DeepOcSortEngine
    from boxmot_b200.synth import make_batch
    from oracle.deepocsort import DeepOcSortOracle
    S, K, F, D, E = 32, 30, 40, 64, 64
    kw = dict(DEEPOCSORT_YAML, aw_off=aw_off)
    dets, nd, embs = make_batch(S, K, F, D, emb_dim=E, first_stream=900)
    eng = DeepOcSortEngine(S, track_cap=128, det_cap=D, emb_dim=E, **kw)
    oracles = [DeepOcSortOracle(**kw) for _ in range(S)]
    rng = np.random.default_rng(5)
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()
    d_embs = torch.as_tensor(embs).cuda()
    for f in range(F):
        w = np.zeros((S, 2, 3))
        w[:, 0, 0] = 1 + rng.normal(0, 0.005, S)
        w[:, 1, 1] = 1 + rng.normal(0, 0.005, S)
        w[:, :, 2] = rng.normal(0, 2, (S, 2))
        out, nout = eng.step(d_dets[f], d_nd[f], d_embs[f], torch.as_tensor(w).cuda())
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
        for s in range(S):
            n = nd[f, s]
            want = oracles[s].update(dets[f, s, :n], None, embs[f, s, :n].copy(), warp=w[s])
            assert_rows_match(out[s, : nout[s]], want, f)
    st = eng.status()
    assert st["track_updates"] == sum(o.n_updates for o in oracles)
    for s in (0, 31):
        x, P, ids = eng.peek(s)
        trk = oracles[s].tracks
        assert np.array_equal(ids, [t.id - 1 for t in trk])
        np.testing.assert_allclose(x, np.array([t.kf.x for t in trk]), rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(P, np.array([t.kf.P for t in trk]), rtol=1e-9, atol=1e-12)


def test_deepocsort_rotation_warps_match_oracle_including_dense_covariance():
    """apply_affine_correction with a similarity warp (rotation): x and y couple, so the engine moves the
    track to its dense 7x7 covariance (kf_xysr.cuh xysr_*) for the rest of its life -- later predicts,
    updates, freezes and ORU replays included.  Rows equal the oracle's; Kalman state within 1e-8."""
    from boxmot_b200.engine import DeepOcSortEngine
    from boxmot_b200.synth import make_batch
    from oracle.deepocsort import DeepOcSortOracle
    S, K, F, D, E = 8, 25, 40, 64, 16
    kw = dict(DEEPOCSORT_YAML, det_thresh=0.4, max_age=6)
    dets, nd, embs = make_batch(S, K, F, D, emb_dim=E, first_stream=1500)
    eng = DeepOcSortEngine(S, track_cap=128, det_cap=D, emb_dim=E, **kw)
    oracles = [DeepOcSortOracle(**kw) for _ in range(S)]
    rng = np.random.default_rng(9)
    d_dets, d_nd, d_embs = torch.as_tensor(dets.astype(np.float32)).cuda(), torch.as_tensor(nd).cuda(), torch.as_tensor(embs).cuda()
    for f in range(F):
        th, sc = rng.normal(0, 0.004, S), 1 + rng.normal(0, 0.005, S)
        if f % 5 == 4:
            th[:] = 0.0                       # rotation-free warps on tracks that are already dense
        w = np.zeros((S, 2, 3))
        w[:, 0, 0] = sc * np.cos(th); w[:, 0, 1] = -sc * np.sin(th)
        w[:, 1, 0] = sc * np.sin(th); w[:, 1, 1] = sc * np.cos(th)
        w[:, :, 2] = rng.normal(0, 2, (S, 2))
        w[S - 1] = [[1, 0, 0], [0, 1, 0]]     # one stream never rotates: stays in block form
        out, nout = eng.step(d_dets[f], d_nd[f], d_embs[f], torch.as_tensor(w).cuda())
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
        for s in range(S):
            n = nd[f, s]
            want = oracles[s].update(dets[f, s, :n], None, embs[f, s, :n].copy(), warp=w[s])
            assert_rows_match(out[s, : nout[s]], want, f)
    eng.status()
    for s in (0, 3, S - 1):
        x, P, ids = eng.peek(s)
        trk = oracles[s].tracks
        assert np.array_equal(ids, [t.id - 1 for t in trk])
        np.testing.assert_allclose(x, np.array([t.kf.x for t in trk]), rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(P, np.array([t.kf.P for t in trk]), rtol=1e-7, atol=1e-9)
        if s < S - 1 and len(trk):
            assert np.abs(P[:, 0, 1]).max() > 0      # the x / y coupling is really there


def test_deepocsort_embedding_decides_between_overlapping_candidates():
    """Two detections swap places between two overlapping tracks: IoU + OCM alone would keep the
    geometric pairing, the embedding term flips it -- the oracle and the engine must agree."""
    from boxmot_b200.engine import DeepOcSortEngine
    from oracle.deepocsort import DeepOcSortOracle
    kw = dict(DEEPOCSORT_YAML, det_thresh=0.3, min_hits=1)
    eng = DeepOcSortEngine(1, track_cap=16, det_cap=16, emb_dim=8, **kw)
    ora = DeepOcSortOracle(**kw)
    ea, eb = np.eye(8, dtype=np.float32)[0], np.eye(8, dtype=np.float32)[1]
    a = np.array([100.0, 100.0, 200.0, 300.0, 0.9, 0.0])
    b = np.array([130.0, 100.0, 230.0, 300.0, 0.9, 0.0])
    seq = [([a, b], [ea, eb])] * 3 + [([a + [4, 0, 4, 0, 0, 0], b - [4, 0, 4, 0, 0, 0]], [eb, ea])] + \
          [([b, a], [ea, eb])] * 2
    for f, (d, e) in enumerate(seq):
        d, e = np.array(d), np.array(e)
        got = eng.step_host([d], [e])[0]
        want = ora.update(d, None, e.copy())
        assert_rows_match(got, want, f)
    eng.status()
