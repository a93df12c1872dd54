"""End-to-end GPU parity of the BoT-SORT engine (tracker of BASELINE configs[2]: cosine
embedding distance + IoU fusion + Kalman XYWH) against reference-generated golden fixtures
and the oracle, through the reference-facing API and the batched engine (C-ABI bm_botsort_*)."""
import numpy as np
import pytest

from conftest import BOTSORT_CASES, BOTSORT_YAML, assert_rows_match, botsort_case, split_by_frame

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
IMG = np.zeros((1080, 1920, 3), np.uint8)


@pytest.mark.parametrize("case", list(BOTSORT_CASES))
def test_botsort_tracker_matches_reference_golden(case):
    """create_tracker("botsort").update(dets, img, embs) row-for-row against the reference:
    ids / conf / cls / det_ind and row order exact, Kalman boxes within 1e-6."""
    import boxmot_b200
    frames, embs, kw, want = botsort_case(case)
    trk = boxmot_b200.create_tracker("botsort", boxmot_b200.get_tracker_config("botsort"),
                                     evolve_param_dict={**BOTSORT_YAML, **kw})
    from conftest import botsort_warps
    warps = botsort_warps(case)
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = trk.update(d, IMG, None if embs is None else embs[f].copy(), warp=None if warps is None else warps[f])
        assert_rows_match(got, want_f[f], f)


def test_botsort_engine_batched_streams_match_oracle():
    """32 streams with 64-d embeddings stepped together; every stream equals a private oracle,
    including Kalman state and the smoothed features' effect on later matches."""
    from boxmot_b200.engine import BotSortEngine
    from boxmot_b200.synth import make_batch
    from oracle.botsort import BotSortOracle
    S, K, F, D, E = 32, 30, 50, 64, 64
    dets, nd, embs = make_batch(S, K, F, D, emb_dim=E, first_stream=500)
    eng = BotSortEngine(S, track_cap=128, det_cap=D, emb_dim=E, **BOTSORT_YAML)
    oracles = [BotSortOracle(**BOTSORT_YAML) for _ in range(S)]
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()
    d_embs = torch.as_tensor(embs).cuda()
    for f in range(F):
        out, nout = eng.step(d_dets[f], d_nd[f], d_embs[f])
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
        for s in range(S):
            n = nd[f, s]
            want = oracles[s].update(dets[f, s, :n], embs[f, s, :n].copy())
            assert_rows_match(out[s, : nout[s]], want, f)
    c = eng.counters()
    assert c["track_updates"] == sum(o.n_updates for o in oracles)
    for s in (0, 31):
        mean, cov, ids, na = eng.peek(s)
        trk = oracles[s].active + oracles[s].lost
        assert np.array_equal(ids, [t.id for t in trk])
        # the block filter reproduces the reference's numpy / scipy arithmetic bit for bit (csrc/kf_xyah.cuh)
        assert np.array_equal(mean, np.array([t.mean for t in trk]))
        assert np.array_equal(cov, np.array([t.cov for t in trk]))


def test_botsort_warps_match_oracle_block_and_dense_covariance():
    """multi_gmc (botsort_track.py:92-104): warps without rotation keep the block covariance and are applied
    exactly; a warp with rotation couples x<->y, w<->h (kron(I4, R)) and moves the track to the dense 8x8
    covariance for the rest of its life (dense predict / Cholesky update of the XYWH filter)."""
    from boxmot_b200.engine import BotSortEngine
    from boxmot_b200.synth import make_stream
    from oracle.botsort import BotSortOracle
    kw = dict(BOTSORT_YAML, with_reid=False)
    eng = BotSortEngine(2, track_cap=64, det_cap=64, **kw)
    ora = [BotSortOracle(**kw), BotSortOracle(**kw)]
    frames = make_stream(21, 20, 45)
    rng = np.random.default_rng(4)
    for f, d in enumerate(frames):
        w0 = np.array([[1.0 + 0.002 * (f % 3), 0.0, 1.5 * (f % 4)], [0.0, 1.0 - 0.001 * (f % 2), -0.75 * (f % 3)]])
        th, sc = rng.normal(0, 0.004), 1 + rng.normal(0, 0.004)
        w1 = np.array([[sc * np.cos(th), -sc * np.sin(th), rng.normal(0, 2)], [sc * np.sin(th), sc * np.cos(th), rng.normal(0, 2)]])
        if f % 4 == 3:
            w1 = w0                              # a rotation-free warp on tracks that are already dense
        got = eng.step_host([d, d], None, [w0, w1])
        for s, w in enumerate((w0, w1)):
            assert_rows_match(got[s], ora[s].update(d, None, w), f)
    eng.status()
    for s in range(2):
        mean, cov, ids, na = eng.peek(s)
        trk = ora[s].active + ora[s].lost
        assert np.array_equal(ids, [t.id for t in trk])
        np.testing.assert_allclose(mean, np.array([t.mean for t in trk]), rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(cov, np.array([t.cov for t in trk]), rtol=1e-7, atol=1e-9)
    assert np.abs(cov[:, 0, 1]).max() > 0            # stream 1 really carries the x / y coupling


def test_botsort_class_votes_follow_the_reference():
    """update_cls (botsort_track.py:51-64): the reported class is the one with the largest
    accumulated confidence, except that a class seen for the first time wins that frame."""
    from boxmot_b200.engine import BotSortEngine
    from oracle.botsort import BotSortOracle
    kw = dict(BOTSORT_YAML, with_reid=False)
    eng = BotSortEngine(1, track_cap=16, det_cap=16, **kw)
    ora = BotSortOracle(**kw)
    cls_seq = [0, 0, 1, 1, 1, 0, 2, 1, 0, 0, 0, 2]
    for f, c in enumerate(cls_seq):
        d = np.array([[100.0 + 2 * f, 200.0, 180.0 + 2 * f, 420.0, 0.9 - 0.01 * f, c],
                      [900.0, 300.0 + f, 960.0, 480.0 + f, 0.8, 3]])
        got = eng.step_host([d])[0]
        want = ora.update(d)
        assert_rows_match(got, want, f)
    eng.status()
