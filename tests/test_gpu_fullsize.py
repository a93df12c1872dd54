"""BASELINE-size runs (1024 concurrent streams) checked through size-independent properties:
streams are independent (a stream's rows do not depend on what else is in the batch), ids are
allocated 1, 2, 3, ... per stream, every row points at a distinct detection of its frame, and a
sample of streams equals the CPU oracle frame by frame."""
import numpy as np
import pytest

from conftest import BYTETRACK_YAML, OCSORT_YAML, assert_rows_match

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
S, K, F, DCAP, TCAP = 1024, 100, 40, 128, 256
SAMPLE = (0, 1, 257, 511, 640, 1023)


def _run(engine_cls, kw, dets, nd):
    eng = engine_cls(dets.shape[1], track_cap=TCAP, det_cap=DCAP, **kw)
    d_dets = torch.as_tensor(np.ascontiguousarray(dets, dtype=np.float32)).cuda()
    d_nd = torch.as_tensor(np.ascontiguousarray(nd)).cuda()
    rows = []
    for f in range(dets.shape[0]):
        out, nout = eng.step(d_dets[f], d_nd[f])
        rows.append((out.cpu().numpy(), nout.cpu().numpy()))
    eng.status()
    return rows, eng.counters()


@pytest.mark.parametrize("tracker", ["ocsort", "bytetrack"])
def test_full_size_batch_properties(tracker):
    from boxmot_b200.engine import ByteTrackEngine, OcSortEngine
    from boxmot_b200.synth import make_batch
    if tracker == "ocsort":
        from oracle.ocsort import OcSortOracle as Oracle
        cls, kw = OcSortEngine, OCSORT_YAML
    else:
        from oracle.bytetrack import ByteTrackOracle as Oracle
        cls, kw = ByteTrackEngine, BYTETRACK_YAML
    dets, nd = make_batch(S, K, F, DCAP)
    rows, cnt = _run(cls, kw, dets, nd)
    # (1) a sample of streams equals the oracle, frame by frame
    oracles = {s: Oracle(**kw) for s in SAMPLE}
    for f in range(F):
        out, nout = rows[f]
        for s in SAMPLE:
            assert_rows_match(out[s, : nout[s]], oracles[s].update(dets[f, s, : nd[f, s]]), f)
    # (2) per stream: ids start at 1 and the set of ids seen so far has no holes larger than the
    #     births that were never confirmed; det_ind are distinct valid rows of the frame; conf / cls
    #     are those of the detection the row points at
    max_id = np.zeros(S)
    for f in range(F):
        out, nout = rows[f]
        for s in range(0, S, 37):
            r = out[s, : nout[s]]
            if not len(r):
                continue
            di = r[:, 7].astype(int)
            assert len(set(di)) == len(di) and di.min() >= 0 and di.max() < nd[f, s]
            assert len(set(r[:, 4])) == len(r) and r[:, 4].min() >= 1
            np.testing.assert_array_equal(r[:, 5], dets[f, s, di, 4].astype(np.float32).astype(np.float64))
            max_id[s] = max(max_id[s], r[:, 4].max())
    assert cnt["births"] >= max_id[::37].sum() > 0
    # (3) independence: the same streams in a different batch (reversed order, 64 of them) give
    #     identical rows
    pick = np.arange(S)[::-16][:64]
    rows2, _ = _run(cls, kw, dets[:, pick], nd[:, pick])
    for f in range(F):
        out, nout = rows[f]
        out2, nout2 = rows2[f]
        for j, s in enumerate(pick):
            assert nout[s] == nout2[j]
            assert np.array_equal(out[s, : nout[s]], out2[j, : nout2[j]]), (f, s)


@pytest.mark.parametrize("tracker", ["deepocsort", "boosttrack"])
def test_full_size_batch_properties_embedding_trackers(tracker):
    """1024 streams with 64-d embeddings through the DeepOCSORT / BoostTrack++ engines: a sample of
    streams equals the oracle frame by frame, rows point at distinct detections of their frame, and
    a stream's rows do not depend on the rest of the batch (bit-identical in another batch)."""
    from boxmot_b200.engine import BoostTrackEngine, DeepOcSortEngine
    from boxmot_b200.synth import make_batch
    from conftest import BOOSTTRACK_YAML, DEEPOCSORT_YAML
    E, FR = 64, 25
    if tracker == "deepocsort":
        from oracle.deepocsort import DeepOcSortOracle as Oracle
        cls, kw, conf_rtol = DeepOcSortEngine, DEEPOCSORT_YAML, 0.0
    else:
        from oracle.boosttrack import BoostTrackOracle as Oracle
        cls, kw, conf_rtol = BoostTrackEngine, BOOSTTRACK_YAML, 1e-9
    dets, nd, embs = make_batch(S, K, FR, DCAP, emb_dim=E)

    def run(dets, nd, embs):
        eng = cls(dets.shape[1], track_cap=TCAP, det_cap=DCAP, emb_dim=E, **kw)
        d_dets = torch.as_tensor(np.ascontiguousarray(dets, dtype=np.float32)).cuda()
        d_nd = torch.as_tensor(np.ascontiguousarray(nd)).cuda()
        d_embs = torch.as_tensor(np.ascontiguousarray(embs)).cuda()
        rows = []
        for f in range(dets.shape[0]):
            out, nout = eng.step(d_dets[f], d_nd[f], d_embs[f])
            rows.append((out.cpu().numpy(), nout.cpu().numpy()))
        eng.status()
        return rows

    rows = run(dets, nd, embs)
    oracles = {s: Oracle(**kw) for s in SAMPLE}
    for f in range(FR):
        out, nout = rows[f]
        for s in SAMPLE:
            n = nd[f, s]
            want = oracles[s].update(dets[f, s, :n], None, embs[f, s, :n].copy())
            assert_rows_match(out[s, : nout[s]], want, f, conf_rtol=conf_rtol)
        for s in range(0, S, 37):
            r = out[s, : nout[s]]
            if len(r):
                di = r[:, 7].astype(int)
                assert len(set(di)) == len(di) and di.min() >= 0 and di.max() < nd[f, s]
                assert len(set(r[:, 4])) == len(r) and r[:, 4].min() >= 1
    pick = np.arange(S)[::-16][:64]
    rows2 = run(dets[:, pick], nd[:, pick], embs[:, pick])
    for f in range(FR):
        out, nout = rows[f]
        out2, nout2 = rows2[f]
        for j, s in enumerate(pick):
            assert nout[s] == nout2[j]
            assert np.array_equal(out[s, : nout[s]], out2[j, : nout2[j]]), (f, s)


@pytest.mark.parametrize("tracker", ["botsort", "strongsort"])
def test_config_shape_batch_properties_botsort_strongsort(tracker):
    """BASELINE configs[2] / configs[3] shapes per stream -- BoT-SORT with 200 detections per frame and 512-d
    embeddings (512 streams), StrongSORT with 100 detections per frame, 768-d embeddings and a 50-deep sample
    bank (96 streams) -- checked through size-independent properties: a sample of streams equals the CPU
    oracle frame by frame, rows point at distinct detections of their frame with ids >= 1, and a stream's
    rows do not depend on the rest of the batch (bit-identical when stepped in another batch)."""
    from boxmot_b200.engine import BotSortEngine, StrongSortEngine
    from boxmot_b200.synth import make_batch
    from conftest import BOTSORT_YAML, STRONGSORT_YAML
    if tracker == "botsort":
        from oracle.botsort import BotSortOracle as Oracle
        NS, KO, FR, E, DC = 512, 200, 10, 512, 256
        kw = BOTSORT_YAML
        make = lambda n: BotSortEngine(n, track_cap=512, det_cap=DC, emb_dim=E, **kw)
        upd = lambda o, d, e: o.update(d, e.copy())
        sample = (0, 255, 511)
    else:
        from oracle.strongsort_tracker import StrongSortOracle as Oracle
        NS, KO, FR, E, DC = 96, 100, 30, 768, 128
        kw = dict(STRONGSORT_YAML, nn_budget=50)
        make = lambda n: StrongSortEngine(n, track_cap=1024, det_cap=DC, pool_cap=4096, emb_dim=E, **kw)
        upd = lambda o, d, e: o.update(d.astype(np.float64), e.astype(np.float64))
        sample = (0, 95)
    dets, nd, embs = make_batch(NS, KO, FR, DC, emb_dim=E)
    dets = dets.astype(np.float32)

    states = {}

    def run(dets, nd, embs, keep=()):
        eng = make(dets.shape[1])
        d_dets, d_nd = torch.as_tensor(np.ascontiguousarray(dets)).cuda(), torch.as_tensor(np.ascontiguousarray(nd)).cuda()
        rows = []
        for f in range(dets.shape[0]):
            out, nout = eng.step(d_dets[f], d_nd[f], torch.as_tensor(np.ascontiguousarray(embs[f])).cuda())
            rows.append((out.cpu().numpy(), nout.cpu().numpy()))
        eng.status()
        for s in keep:
            states[s] = eng.peek(s)
        eng.close()
        return rows

    rows = run(dets, nd, embs, keep=sample)
    oracles = {s: Oracle(**kw) for s in sample}
    for f in range(FR):
        out, nout = rows[f]
        for s in sample:
            n = nd[f, s]
            assert_rows_match(out[s, : nout[s]], upd(oracles[s], dets[f, s, :n], embs[f, s, :n]), f)
        for s in range(0, NS, 7):
            r = out[s, : nout[s]]
            if len(r):
                di = r[:, 7].astype(int)
                assert len(set(di)) == len(di) and di.min() >= 0 and di.max() < nd[f, s]
                assert len(set(r[:, 4])) == len(r) and r[:, 4].min() >= 1
    # Kalman state of the sampled streams after the last frame: BIT-IDENTICAL to the oracle's numpy / scipy
    # filter (csrc/kf_xyah.cuh carries both off-diagonal covariance entries and scipy's reciprocal solves)
    for s in sample:
        o = oracles[s]
        trk = o.tracks if tracker == "strongsort" else o.active + o.lost
        assert np.array_equal(states[s][0], np.array([t.mean for t in trk]))
        assert np.array_equal(states[s][1], np.array([t.cov for t in trk]))
    pick = np.arange(NS)[::-5][:16]
    rows2 = run(dets[:, pick], nd[:, pick], embs[:, pick])
    for f in range(FR):
        out, nout = rows[f]
        out2, nout2 = rows2[f]
        for j, s in enumerate(pick):
            assert nout[s] == nout2[j]
            assert np.array_equal(out[s, : nout[s]], out2[j, : nout2[j]]), (f, s)


@pytest.mark.parametrize("tracker", ["deepocsort", "boosttrack"])
def test_crowd_density_streams_match_oracle(tracker):
    """BASELINE configs[4] shape per stream: MOT20-like crowds of ~480 detections per frame (hot-spot
    clustered, heavy mutual overlap) with 64-d embeddings; every stream equals a private oracle frame
    by frame.  Caps: 768 tracks x 512 detections per stream (BoostTrack's per-stream tables then fill
    215 KB of the 227 KB of shared memory)."""
    from boxmot_b200.engine import BoostTrackEngine, DeepOcSortEngine
    from boxmot_b200.synth import make_batch
    from conftest import BOOSTTRACK_YAML, DEEPOCSORT_YAML
    NS, KC, FR, E, TC, DC = 6, 480, 14, 64, 768, 512
    if tracker == "deepocsort":
        from oracle.deepocsort import DeepOcSortOracle as Oracle
        cls, kw, conf_rtol = DeepOcSortEngine, DEEPOCSORT_YAML, 0.0
    else:
        from oracle.boosttrack import BoostTrackOracle as Oracle
        cls, kw, conf_rtol = BoostTrackEngine, BOOSTTRACK_YAML, 1e-9
    dets, nd, embs = make_batch(NS, KC, FR, DC, emb_dim=E, crowd=True, first_stream=2000)
    assert nd.max() > 470
    eng = cls(NS, track_cap=TC, det_cap=DC, emb_dim=E, **kw)
    oracles = [Oracle(**kw) for _ in range(NS)]
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()
    d_embs = torch.as_tensor(embs).cuda()
    total = 0
    for f in range(FR):
        out, nout = eng.step(d_dets[f], d_nd[f], d_embs[f])
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
        for s in range(NS):
            n = nd[f, s]
            want = oracles[s].update(dets[f, s, :n], None, embs[f, s, :n].copy())
            assert_rows_match(out[s, : nout[s]], want, f, conf_rtol=conf_rtol)
            total += len(want)
    eng.status()
    assert total > 5000


@pytest.mark.parametrize("tracker", ["ocsort", "deepocsort", "boosttrack", "botsort"])
def test_stream_groups_do_not_change_results(tracker):
    """bm_ocsort_set_groups: the same batch stepped as 1, 3 and 8 groups of streams on CUDA streams of
    their own gives bit-identical rows and counters (streams are independent; 250 is not a multiple
    of 3 or 8, so the groups are ragged)."""
    from boxmot_b200.engine import BoostTrackEngine, BotSortEngine, DeepOcSortEngine, OcSortEngine
    from boxmot_b200.synth import make_batch
    from conftest import BOOSTTRACK_YAML, BOTSORT_YAML, DEEPOCSORT_YAML, OCSORT_YAML
    NS, KO, FR, E = 250, 60, 20, 32
    if tracker == "ocsort":
        dets, nd = make_batch(NS, KO, FR, 96)
        embs = None
    else:
        dets, nd, embs = make_batch(NS, KO, FR, 96, emb_dim=E)
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()
    d_embs = None if embs is None else torch.as_tensor(embs).cuda()
    rng = np.random.default_rng(3)
    warps = np.zeros((FR, NS, 2, 3))
    warps[:, :, 0, 0] = 1 + rng.normal(0, 0.004, (FR, NS))
    warps[:, :, 1, 1] = 1 + rng.normal(0, 0.004, (FR, NS))
    warps[:, :, :, 2] = rng.normal(0, 2, (FR, NS, 2))
    d_w = torch.as_tensor(warps).cuda()

    def run(groups):
        if tracker == "ocsort":
            eng = OcSortEngine(NS, track_cap=192, det_cap=96, **OCSORT_YAML)
        elif tracker == "deepocsort":
            eng = DeepOcSortEngine(NS, track_cap=192, det_cap=96, emb_dim=E, **DEEPOCSORT_YAML)
        elif tracker == "botsort":
            eng = BotSortEngine(NS, track_cap=192, det_cap=96, emb_dim=E, **BOTSORT_YAML)
        else:
            eng = BoostTrackEngine(NS, track_cap=192, det_cap=96, emb_dim=E, **BOOSTTRACK_YAML)
        assert eng.set_stream_groups(groups) == groups
        rows = []
        for f in range(FR):
            out, nout = eng.step(d_dets[f], d_nd[f]) if tracker == "ocsort" else \
                eng.step(d_dets[f], d_nd[f], d_embs[f], d_w[f])
            rows.append((out.cpu().numpy().copy(), nout.cpu().numpy().copy()))
        return rows, eng.status()

    base, st0 = run(1)
    assert st0["track_updates"] > 0
    for g in (3, 8):
        rows, st = run(g)
        assert st == st0
        for f in range(FR):
            assert np.array_equal(rows[f][1], base[f][1])
            for s in range(NS):
                n = base[f][1][s]
                assert np.array_equal(rows[f][0][s, :n], base[f][0][s, :n]), (g, f, s)


def test_prescreen_launch_does_not_change_results(monkeypatch):
    """ocs_screen_kernel (the first-round fp32 screen as its own launch, default for batches of more
    than 1536 streams; BM_PRESCREEN forces it): rows and counters bit-identical to the in-kernel
    screen, and a sample of streams equals the oracle."""
    from boxmot_b200.engine import OcSortEngine
    from boxmot_b200.synth import make_batch
    from conftest import OCSORT_YAML
    from oracle.ocsort import OcSortOracle
    NS, KO, FR = 96, 100, 30
    dets, nd = make_batch(NS, KO, FR, 128, first_stream=300)
    d_dets = torch.as_tensor(dets.astype(np.float32)).cuda()
    d_nd = torch.as_tensor(nd).cuda()

    def run(flag):
        monkeypatch.setenv("BM_PRESCREEN", flag)
        eng = OcSortEngine(NS, track_cap=256, det_cap=128, **OCSORT_YAML)
        rows = []
        for f in range(FR):
            out, nout = eng.step(d_dets[f], d_nd[f])
            rows.append((out.cpu().numpy().copy(), nout.cpu().numpy().copy()))
        return rows, eng.status()

    base, st0 = run("0")
    rows, st1 = run("1")
    assert st0 == st1 and st0["track_updates"] > 0
    oracles = {s: OcSortOracle(**OCSORT_YAML) for s in (0, 41, 95)}
    for f in range(FR):
        assert np.array_equal(rows[f][1], base[f][1])
        for s in range(NS):
            n = base[f][1][s]
            assert np.array_equal(rows[f][0][s, :n], base[f][0][s, :n]), (f, s)
        for s, o in oracles.items():
            want = o.update(dets[f, s, : nd[f, s]])
            assert_rows_match(rows[f][0][s, : rows[f][1][s]], want, f)
