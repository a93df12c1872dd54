"""End-to-end GPU parity of the ByteTrack engine (BASELINE configs[0]: MOT17-mini detections,
bytetrack.yaml) against the reference-generated golden fixtures and the oracle, through the
reference-facing API and the batched engine (C-ABI bm_bytetrack_*)."""
import numpy as np
import pytest

from conftest import (BYTETRACK_YAML, SYNTH_CASES, assert_rows_match, case_frames, golden_rows,
                      split_by_frame)

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
IMG = np.zeros((1080, 1920, 3), np.uint8)


@pytest.mark.parametrize("case", list(SYNTH_CASES) + ["MOT17_02_FRCNN", "MOT17_04_FRCNN"])
def test_bytetrack_tracker_matches_reference_golden(case):
    """create_tracker("bytetrack").update(dets, img) row-for-row against the reference's output:
    ids / conf / cls / det_ind and row order exact, Kalman boxes within 1e-6."""
    import boxmot_b200
    want, _ = golden_rows("bytetrack", case)
    frames = case_frames(case)
    trk = boxmot_b200.create_tracker("bytetrack", boxmot_b200.get_tracker_config("bytetrack"))
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = trk.update(d, IMG)
        assert_rows_match(got, want_f[f], f)


def test_bytetrack_engine_batched_streams_match_oracle():
    """48 different streams stepped by one launch per frame; every stream equals a private
    oracle instance, including the Kalman state of every active and lost track."""
    from boxmot_b200.engine import ByteTrackEngine
    from boxmot_b200.synth import make_batch
    from oracle.bytetrack import ByteTrackOracle
    S, K, F, D = 48, 40, 60, 64
    dets, nd = make_batch(S, K, F, D, first_stream=300)
    eng = ByteTrackEngine(S, track_cap=128, det_cap=D, **BYTETRACK_YAML)
    oracles = [ByteTrackOracle(**BYTETRACK_YAML) for _ in range(S)]
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()
    for f in range(F):
        out, nout = eng.step(d_dets[f], d_nd[f])
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
        for s in range(S):
            want = oracles[s].update(dets[f, s, : nd[f, s]])
            assert_rows_match(out[s, : nout[s]], want, f)
    c = eng.counters()
    assert c["track_updates"] == sum(o.n_updates for o in oracles)
    for s in (0, 17, 47):
        mean, cov, ids, na = eng.peek(s)
        o = oracles[s]
        trk = o.active + o.lost
        assert na == len(o.active) and len(ids) == len(trk)
        assert np.array_equal(ids, [t.id for t in trk])
        # the block filter reproduces the reference's numpy / scipy arithmetic bit for bit (csrc/kf_xyah.cuh)
        assert np.array_equal(mean, np.array([t.mean for t in trk]))
        assert np.array_equal(cov, np.array([t.cov for t in trk]))


def test_bytetrack_lost_tracks_are_refound_and_age_out():
    """Objects that disappear for a while: lost -> re-found keeps the id; lost for more than
    track_buffer frames -> removed one frame late, exactly as the reference's list handling."""
    from boxmot_b200.engine import ByteTrackEngine
    from oracle.bytetrack import ByteTrackOracle
    kw = dict(BYTETRACK_YAML, track_buffer=5)
    eng = ByteTrackEngine(1, track_cap=64, det_cap=32, **kw)
    ora = ByteTrackOracle(**kw)
    rng = np.random.default_rng(5)
    c0 = rng.random((8, 2)) * [1500, 800] + 150
    vel = rng.normal(0, 2.0, (8, 2))
    for f in range(60):
        vis = np.ones(8, bool)
        vis[0] = not (10 <= f < 13)            # short gap: re-found
        vis[1] = not (15 <= f < 30)            # long gap: aged out, then a new id
        vis[2] = f % 7 != 3                    # flicker
        vis[3] = f < 20 or f >= 26             # gap == track_buffer + 1 (the one-frame-late removal)
        c = c0 + vel * f
        wh = np.array([60.0, 140.0])
        d = np.c_[c - wh / 2, c + wh / 2, np.full(8, 0.9), np.zeros(8)][vis]
        d[:, :4] += rng.normal(0, 1.0, (len(d), 4))
        if f % 5 == 0 and len(d):
            d[0, 4] = 0.3                      # low score: second association
        got = eng.step_host([d])[0]
        want = ora.update(d)
        assert_rows_match(got, want, f)
    eng.status()


def test_bytetrack_empty_frames_and_capacity_error():
    from boxmot_b200.engine import ByteTrackEngine
    eng = ByteTrackEngine(1, track_cap=8, det_cap=32, **BYTETRACK_YAML)
    assert len(eng.step_host([np.empty((0, 6))])[0]) == 0
    rng = np.random.default_rng(1)
    c = rng.random((20, 2)) * [1600, 900] + 100
    d = np.c_[c, c + [50, 120], np.full(20, 0.9), np.zeros(20)]
    eng.step_host([d])
    with pytest.raises(RuntimeError, match="capacity"):
        eng.status()
    with pytest.raises(RuntimeError, match="det_cap"):
        eng.step_host([np.zeros((40, 6))])


def test_bytetrack_input_validation_matches_reference_asserts():
    import boxmot_b200
    trk = boxmot_b200.create_tracker("bytetrack", boxmot_b200.get_tracker_config("bytetrack"))
    with pytest.raises(AssertionError):
        trk.update(np.zeros((3, 5)), IMG)
    with pytest.raises(AssertionError):
        trk.update([[0, 0, 1, 1, 0.9, 0]], IMG)
    out = trk.update(np.empty((0, 6)), IMG)
    assert out.size == 0
