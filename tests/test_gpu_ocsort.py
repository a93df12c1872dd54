"""End-to-end GPU parity of the OC-SORT engine against the reference-generated golden
fixtures and the oracle, through the reference-facing API and the batched engine."""
import numpy as np
import pytest

from conftest import (OCSORT_YAML, SYNTH_CASES, assert_rows_match, case_frames, golden_rows,
                      split_by_frame)
from boxmot_b200.synth import make_stream as case_frames_raw

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
IMG = np.zeros((1080, 1920, 3), np.uint8)


@pytest.mark.parametrize("case", list(SYNTH_CASES) + ["MOT17_02_FRCNN", "MOT17_04_FRCNN"])
def test_ocsort_tracker_matches_reference_golden(case):
    """create_tracker(...).update(dets, img) row-for-row against the reference's output."""
    import boxmot_b200
    want, kw = golden_rows("ocsort", case)
    frames = case_frames(case)
    trk = boxmot_b200.create_tracker("ocsort", boxmot_b200.get_tracker_config("ocsort"),
                                     evolve_param_dict={**OCSORT_YAML, **kw})
    want_f = split_by_frame(want, len(frames))
    for f, d in enumerate(frames):
        got = trk.update(d, IMG)
        assert_rows_match(got, want_f[f], f)


def test_ocsort_engine_batched_streams_match_oracle():
    """64 different streams stepped by one launch per frame; every stream must equal a
    private oracle instance (ids are per stream)."""
    from boxmot_b200.engine import OcSortEngine
    from boxmot_b200.synth import make_batch
    from oracle.ocsort import OcSortOracle
    S, K, F, D = 64, 40, 50, 64
    dets, nd = make_batch(S, K, F, D, first_stream=100)
    eng = OcSortEngine(S, track_cap=128, det_cap=D, **OCSORT_YAML)
    oracles = [OcSortOracle(**OCSORT_YAML) for _ in range(S)]
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()
    for f in range(F):
        out, nout = eng.step(d_dets[f], d_nd[f])
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
        for s in range(S):
            want = oracles[s].update(dets[f, s, : nd[f, s]])
            assert_rows_match(out[s, : nout[s]], want, f)
    st = eng.status()
    assert st["track_updates"] == sum(o.n_updates for o in oracles)


def test_ocsort_empty_frames_and_capacity_error():
    from boxmot_b200.engine import OcSortEngine
    from oracle.ocsort import OcSortOracle
    eng = OcSortEngine(1, track_cap=32, det_cap=64, **OCSORT_YAML)
    ora = OcSortOracle(**OCSORT_YAML)
    rng = np.random.default_rng(0)
    base = rng.random((10, 2)) * [1500, 800] + 100
    for f in range(12):
        if f in (3, 4, 9):
            d = np.empty((0, 6))
        else:
            c = base + 2 * f
            d = np.c_[c, c + [50, 110], rng.uniform(0.65, 0.95, 10), np.zeros(10)]
        got = eng.step_host([d])[0]
        assert_rows_match(got, ora.update(d), f)
    eng.status()
    # 40 far-apart detections into a 32-track engine must raise, not truncate
    c = np.stack([np.arange(40) * 45.0 + 10, np.full(40, 100.0)], 1)
    d = np.c_[c, c + [30, 60], np.full(40, 0.9), np.zeros(40)]
    eng2 = OcSortEngine(1, track_cap=32, det_cap=64, **OCSORT_YAML)
    eng2.step_host([d])
    with pytest.raises(RuntimeError):
        eng2.status()


def test_input_validation_matches_reference_asserts():
    import boxmot_b200
    trk = boxmot_b200.create_tracker("ocsort", boxmot_b200.get_tracker_config("ocsort"))
    with pytest.raises(AssertionError):
        trk.update(np.zeros((3, 5)), IMG)
    with pytest.raises(AssertionError):
        trk.update([[0, 0, 1, 1, 0.9, 0]], IMG)
    out = trk.update(np.empty((0, 6)), IMG)
    assert out.size == 0


def test_host_pipeline_matches_direct_stepping():
    """The double-buffered H2D / step / pack-into-pinned-host-memory pipeline returns, one frame late,
    exactly the rows a synchronous step returns, as one dense ragged buffer (rows, offsets)."""
    from boxmot_b200.engine import HostPipeline, OcSortEngine
    from boxmot_b200.synth import make_batch
    S, K, F, D = 16, 30, 20, 64
    dets, nd = make_batch(S, K, F, D, first_stream=300)
    h_dets = torch.as_tensor(dets.astype(np.float32)).pin_memory()
    h_nd = torch.as_tensor(nd).pin_memory()
    ref = OcSortEngine(S, track_cap=128, det_cap=D, **OCSORT_YAML)
    want = []
    for f in range(F):
        out, nout = ref.step(h_dets[f].cuda(), h_nd[f].cuda())
        want.append((out.cpu().numpy().copy(), nout.cpu().numpy().copy()))
    eng = OcSortEngine(S, track_cap=128, det_cap=D, **OCSORT_YAML)
    pipe = HostPipeline(eng)
    got = []
    for f in range(F):
        r = pipe.submit(h_dets[f], h_nd[f])
        if r is not None:
            got.append((r[0].copy(), r[1].copy()))
    r = pipe.drain()
    got.append((r[0].copy(), r[1].copy()))
    eng.status()
    assert len(got) == F
    for f in range(F):
        rows, off = got[f]
        assert np.array_equal(np.diff(off), want[f][1]) and off[0] == 0 and len(rows) == off[-1]
        for s in range(S):
            n = want[f][1][s]
            assert np.array_equal(rows[off[s]:off[s + 1]], want[f][0][s, :n])


def test_host_pipeline_with_embeddings_and_warps_matches_direct_stepping():
    """The same pipeline for an engine with embedding and warp side inputs (DeepOCSORT)."""
    from conftest import DEEPOCSORT_YAML
    from boxmot_b200.engine import DeepOcSortEngine, HostPipeline
    from boxmot_b200.synth import make_batch
    S, K, F, D, E = 8, 25, 15, 64, 32
    dets, nd, embs = make_batch(S, K, F, D, emb_dim=E, first_stream=340)
    h_dets = torch.as_tensor(dets.astype(np.float32)).pin_memory()
    h_nd = torch.as_tensor(nd).pin_memory()
    h_embs = torch.as_tensor(embs).pin_memory()
    rng = np.random.default_rng(2)
    w = np.zeros((F, S, 2, 3))
    w[:, :, 0, 0] = 1 + rng.normal(0, 0.01, (F, S)); w[:, :, 1, 1] = 1 + rng.normal(0, 0.01, (F, S))
    w[:, :, :, 2] = rng.normal(0, 2, (F, S, 2))
    h_w = torch.as_tensor(w).pin_memory()
    ref = DeepOcSortEngine(S, track_cap=128, det_cap=D, emb_dim=E, **DEEPOCSORT_YAML)
    want = []
    for f in range(F):
        out, nout = ref.step(h_dets[f].cuda(), h_nd[f].cuda(), h_embs[f].cuda(), h_w[f].cuda())
        want.append((out.cpu().numpy().copy(), nout.cpu().numpy().copy()))
    eng = DeepOcSortEngine(S, track_cap=128, det_cap=D, emb_dim=E, **DEEPOCSORT_YAML)
    pipe = HostPipeline(eng)
    got = []
    for f in range(F):
        r = pipe.submit(h_dets[f], h_nd[f], h_embs[f], h_w[f])
        if r is not None:
            got.append((r[0].copy(), r[1].copy()))
    r = pipe.drain()
    got.append((r[0].copy(), r[1].copy()))
    eng.status()
    for f in range(F):
        rows, off = got[f]
        assert np.array_equal(np.diff(off), want[f][1])
        for s in range(S):
            assert np.array_equal(rows[off[s]:off[s + 1]], want[f][0][s, : want[f][1][s]])


def test_ocsort_per_class_matches_reference_golden():
    """per_class=True (basetracker.py:154-201): one list of tracks per class id, ONE id counter
    for the whole tracker, det_ind relative to the class subset, rows stacked in class order."""
    import boxmot_b200
    want, _ = golden_rows("ocsort", "synth_k40_perclass")
    frames = []
    for d in case_frames_raw(5, 40, 60):
        d = d.copy()
        d[:, 5] = np.floor((d[:, 0] + d[:, 2]) / 2.0 / 640.0).clip(0, 2)
        frames.append(d)
    trk = boxmot_b200.create_tracker("ocsort", boxmot_b200.get_tracker_config("ocsort"), per_class=True)
    want_f = split_by_frame(want, len(frames))
    seen = 0
    for f, d in enumerate(frames):
        got = trk.update(d, IMG)
        assert got.shape[1:] == (8,) or got.size == 0
        assert_rows_match(got, want_f[f], f)
        seen += len(got)
    assert seen == len(want) and len(np.unique(want[:, 7])) == 3     # rows of three classes


def test_two_live_engines_with_different_capacities_step_alternately():
    """The dynamic shared-memory opt-in of a kernel is process-global: an engine created with small
    capacities must not lower it under a live engine with large ones (and vice versa).  OC-SORT, ByteTrack
    and BoostTrack engines of two sizes each are stepped alternately; results equal separately stepped engines."""
    from conftest import BOOSTTRACK_YAML, BYTETRACK_YAML
    from boxmot_b200.engine import BoostTrackEngine, ByteTrackEngine, OcSortEngine
    from boxmot_b200.synth import make_batch
    S, K, F = 4, 30, 12
    for make in (lambda t, d: OcSortEngine(S, track_cap=t, det_cap=d, **OCSORT_YAML),
                 lambda t, d: ByteTrackEngine(S, track_cap=t, det_cap=d, **BYTETRACK_YAML),
                 lambda t, d: BoostTrackEngine(S, track_cap=t, det_cap=d, **dict(BOOSTTRACK_YAML, with_reid=False))):
        outs = {}
        for order in ("solo", "mixed"):
            big = make(512, 512)
            small = make(128, 64) if order == "mixed" else None
            got = []
            for caps, eng in ((512, big), (64, small)):
                if eng is None:
                    continue
                dets, nd = make_batch(S, K, F, caps, first_stream=900)
                eng._d = torch.as_tensor(dets.astype(np.float32)).cuda()
                eng._n = torch.as_tensor(nd).cuda()
            for f in range(F):
                o, n = big.step(big._d[f], big._n[f])
                got.append((o.cpu().numpy().copy(), n.cpu().numpy().copy()))
                if small is not None:
                    small.step(small._d[f], small._n[f])
            big.status()
            if small is not None:
                small.status()
                small.close()
            big.close()
            outs[order] = got
        for (a, na), (b, nb) in zip(outs["solo"], outs["mixed"]):
            assert np.array_equal(na, nb)
            for s in range(S):
                assert np.array_equal(a[s, : na[s]], b[s, : nb[s]])
