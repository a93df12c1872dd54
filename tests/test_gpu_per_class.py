"""per_class=True for the trackers whose classes the reference's decorator COUPLES (row a25): ByteTrack and
BoT-SORT share lost_stracks / removed_stracks across the class calls (only active_tracks is swapped,
basetracker.py:183-192), BoostTrack's self.trackers is not swapped at all (boosttrack.py:185).  The fixtures
are the reference's own output in that mode (oracle/gen_golden.py --per-class-shared): 40 objects in three
classes by image column become 311 ByteTrack ids, because every class call also predicts and re-finds the
other classes' lost tracks.  The adapters reproduce it through bm_bytetrack_select_class /
bm_boosttrack_hold_frame: nr_classes engine steps per frame on one stream, no host synchronisation between."""
import numpy as np
import pytest

from conftest import assert_rows_match, deepocsort_per_class_case, golden_rows, split_by_frame

pytestmark = pytest.mark.gpu
IMG = np.zeros((1080, 1920, 3), np.uint8)


def _make(name, **over):
    """create_tracker with the YAML defaults, overridden the way the reference's tuner does (evolve_param_dict)."""
    import yaml
    import boxmot_b200
    over.pop("per_class", None)
    with open(boxmot_b200.get_tracker_config(name)) as f:
        args = {k: v["default"] for k, v in yaml.safe_load(f).items()}
    args.update(over)
    return boxmot_b200.create_tracker(name, evolve_param_dict=args, per_class=True)


def _frames_plain():
    from boxmot_b200.synth import make_stream
    out = []
    for d in make_stream(5, 40, 60):
        d = d.copy()
        d[:, 5] = np.floor((d[:, 0] + d[:, 2]) / 2.0 / 640.0).clip(0, 2)
        out.append(d)
    return out


def _run(trk, frames, embs, want):
    want_f = split_by_frame(want, len(frames))
    seen = 0
    for f, d in enumerate(frames):
        got = trk.update(d, IMG, None if embs is None else embs[f])
        got = np.asarray(got, dtype=float).reshape(-1, 8)
        assert_rows_match(got, want_f[f], f)
        seen += len(got)
    assert seen == len(want)


def test_bytetrack_per_class_matches_reference_golden():
    want, _ = golden_rows("bytetrack", "synth_k40_perclass")
    trk = _make("bytetrack")
    _run(trk, _frames_plain(), None, want)
    assert want[:, 5].max() > 250 and len(np.unique(want[:, 7])) == 3      # the coupling's id churn is in the fixture


def test_botsort_per_class_matches_reference_golden():
    want, _ = golden_rows("botsort", "synth_k40_e32_perclass")
    frames, embs, _, _ = deepocsort_per_class_case()
    trk = _make("botsort")
    _run(trk, frames, embs, want)
    assert want[:, 5].max() > 200


@pytest.mark.parametrize("case", ["synth_k40_perclass", "synth_k40_perclass_age200"])
def test_boosttrack_per_class_matches_reference_golden(case):
    want, kw = golden_rows("boosttrack", case)
    frames, embs, _, _ = deepocsort_per_class_case()
    trk = _make("boosttrack", **kw)
    _run(trk, frames, embs if trk.with_reid else None, want)


def test_per_class_reset_starts_over():
    """reset() clears the per-class lists too: the same frames give the same rows again."""
    trk = _make("bytetrack")
    frames = _frames_plain()[:15]
    a = [np.asarray(trk.update(d, IMG), dtype=float).reshape(-1, 8) for d in frames]
    trk.reset()
    b = [np.asarray(trk.update(d, IMG), dtype=float).reshape(-1, 8) for d in frames]
    for x, y in zip(a, b):
        # ids restart from 1 after reset (the engine's counter), everything else is identical
        assert x.shape == y.shape and np.array_equal(x[:, :4], y[:, :4]) and np.array_equal(x[:, 5:], y[:, 5:])
