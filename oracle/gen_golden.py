"""Generate the committed fixtures under tests/golden/ by running the UNMODIFIED
reference (through oracle/ref_shim.py) in the authoring container.

    python -m oracle.gen_golden

Outputs (npz, compressed):
  mot17_mini_dets.npz      MOT17-mini det.txt of both sequences as [frame, x1,y1,x2,y2,conf]
                           (assets/MOT17-mini/train/*/det/det.txt, converted from ltwh)
  ref_<tracker>_<case>.npz reference outputs: rows [frame, x1,y1,x2,y2,id,conf,cls,det_ind]
The GPU box has no /root/reference; tests read only these files."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402
from boxmot_b200.synth import make_stream, mot17_embs, strongsort_stream  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SEQS = ("MOT17-02-FRCNN", "MOT17-04-FRCNN")
IMG = np.zeros((1080, 1920, 3), np.uint8)


def load_det_txt(seq):
    a = np.loadtxt(os.path.join(ref_shim.REF_ROOT, "assets", "MOT17-mini", "train", seq, "det", "det.txt"),
                   delimiter=",")
    rows = np.c_[a[:, 0], a[:, 2], a[:, 3], a[:, 2] + a[:, 4], a[:, 3] + a[:, 5], a[:, 6]]
    return rows[np.argsort(rows[:, 0], kind="stable")]


def frames_of(rows):
    n = int(rows[:, 0].max())
    return [np.c_[rows[rows[:, 0] == f][:, 1:], np.zeros((rows[:, 0] == f).sum())] for f in range(1, n + 1)]


def run_ref(name, frames, embs=None, **kw):
    trk = ref_shim.make_tracker(name, **kw)
    out = []
    for f, d in enumerate(frames):
        r = trk.update(d.copy(), IMG, None if embs is None else embs[f].copy())
        r = np.asarray(r, dtype=float)
        r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
        out.append(np.c_[np.full(len(r), f), r])
    return np.vstack(out)


# (case name, frames factory, tracker kwargs)
def synth_cases():
    return [
        ("synth_k100", lambda: make_stream(0, 100, 80), {}),
        ("synth_k30_byte", lambda: make_stream(1, 30, 120), {"use_byte": True}),
        ("synth_k30_giou", lambda: make_stream(2, 30, 80), {"asso_func": "giou"}),
        ("synth_k50_dt1", lambda: make_stream(3, 50, 80), {"delta_t": 1, "det_thresh": 0.3}),
        ("synth_k40_ciou", lambda: make_stream(4, 40, 60), {"asso_func": "ciou", "max_age": 5}),
    ]


def per_class_frames():
    """Three classes by image column (class = floor(xc / 640)): objects that cross a boundary
    change class, so tracks die and are born in another class's list."""
    out = []
    for d in make_stream(5, 40, 60):
        d = d.copy()
        d[:, 5] = np.floor((d[:, 0] + d[:, 2]) / 2.0 / 640.0).clip(0, 2)
        out.append(d)
    return out


def main_per_class():
    out = run_ref("ocsort", per_class_frames(), per_class=True)
    np.savez_compressed(os.path.join(GOLD, "ref_ocsort_synth_k40_perclass.npz"), rows=out,
                        kw=np.array(repr({"per_class": True})))
    print("ocsort per_class", out.shape, "ids", int(out[:, 5].max()))


def per_class_emb_frames():
    """per_class_frames() with 32-d embeddings (stream 26)."""
    fr, em = make_stream(26, 40, 60, emb_dim=32)
    out = []
    for d in fr:
        d = d.copy()
        d[:, 5] = np.floor((d[:, 0] + d[:, 2]) / 2.0 / 640.0).clip(0, 2)
        out.append(d)
    return out, em


def main_deepocsort_per_class():
    """DeepOCSORT with per_class=True: the decorator (basetracker.py:155-201) calls update once per
    class id, so the CMC stand-in is asked nr_classes times per frame -- queue the frame's warp
    that many times."""
    fr, em = per_class_emb_frames()
    warps = deepocsort_warps(26, len(fr))
    trk = ref_shim.make_tracker("deepocsort", per_class=True)
    out = []
    for i, d in enumerate(fr):
        trk.cmc.warps = [warps[i].copy() for _ in range(trk.nr_classes)]
        r = np.asarray(trk.update(d.copy(), IMG, em[i].copy()), dtype=float)
        r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
        out.append(np.c_[np.full(len(r), i), r])
    out = np.vstack(out)
    np.savez_compressed(os.path.join(GOLD, "ref_deepocsort_synth_k40_perclass.npz"), rows=out,
                        kw=np.array(repr({"per_class": True})), warps=warps)
    print("deepocsort per_class", out.shape, "ids", int(out[:, 5].max()))


def main_per_class_shared():
    """per_class=True for the trackers whose classes are COUPLED by the decorator (basetracker.py:154-201):
    ByteTrack / BoT-SORT share lost_stracks and removed_stracks across the class calls, BoostTrack's self.trackers
    is not swapped at all.  The CMC stand-in is asked nr_classes times per frame: identity every time."""
    out = run_ref("bytetrack", per_class_frames(), per_class=True)
    np.savez_compressed(os.path.join(GOLD, "ref_bytetrack_synth_k40_perclass.npz"), rows=out,
                        kw=np.array(repr({"per_class": True})))
    print("bytetrack per_class", out.shape, "ids", int(out[:, 5].max()))
    fr, em = per_class_emb_frames()
    for name, case, kw in (("botsort", "synth_k40_e32_perclass", {}),
                           ("boosttrack", "synth_k40_perclass", {}),
                           ("boosttrack", "synth_k40_perclass_age200", {"max_age": 200, "with_reid": True})):
        trk = ref_shim.make_tracker(name, per_class=True, **kw)
        rows = []
        for i, d in enumerate(fr):
            r = np.asarray(trk.update(d.copy(), IMG, em[i].copy()), dtype=float)
            r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
            rows.append(np.c_[np.full(len(r), i), r])
        out = np.vstack(rows)
        np.savez_compressed(os.path.join(GOLD, "ref_%s_%s.npz" % (name, case)), rows=out,
                            kw=np.array(repr(dict(kw, per_class=True))))
        print(name, case, out.shape, "ids", int(out[:, 5].max()) if len(out) else None)


def main_kf():
    """Known-answer vectors of the reference's KalmanFilterXYAH / XYWH (rows a1-a5): a chain
    initiate -> (predict -> update(conf))* on random boxes, and gating distances."""
    ref_shim.load()
    from boxmot.motion.kalman_filters.aabb.xyah_kf import KalmanFilterXYAH
    from boxmot.motion.kalman_filters.aabb.xywh_kf import KalmanFilterXYWH
    rng = np.random.default_rng(11)
    out = {}
    for kind, kf in (("xyah", KalmanFilterXYAH()), ("xywh", KalmanFilterXYWH())):
        n, steps = 64, 6
        c = rng.random((n, 2)) * [1800, 1000]
        wh = rng.uniform(20, 300, (n, 2))
        z0 = np.c_[c, wh[:, 0] / wh[:, 1] if kind == "xyah" else wh[:, 0], wh[:, 1]]
        zs = z0[None] + np.cumsum(rng.normal(0, 1.5, (steps, n, 4)) * [1, 1, 0.01 if kind == "xyah" else 1, 1], axis=0)
        conf = rng.uniform(0, 0.95, (steps, n))
        means, covs = [], []
        cur = [kf.initiate(z) for z in z0]
        means.append(np.array([m for m, _ in cur])); covs.append(np.array([c_ for _, c_ in cur]))
        for k in range(steps):
            cur = [kf.predict(m, c_) for m, c_ in cur]
            means.append(np.array([m for m, _ in cur])); covs.append(np.array([c_ for _, c_ in cur]))
            cur = [kf.update(m, c_, zs[k, i], conf[k, i] if k % 2 else 0.0) for i, (m, c_) in enumerate(cur)]
            means.append(np.array([m for m, _ in cur])); covs.append(np.array([c_ for _, c_ in cur]))
        mm, cc = kf.multi_predict(means[-1].copy(), covs[-1].copy())
        meas = zs[-1][:20]
        gate = {}
        for op in (False, True):
            for metric in ("maha", "gaussian"):
                gate["%s_%d" % (metric, op)] = np.array([kf.gating_distance(m, c_, meas, op, metric)
                                                         for m, c_ in cur])
        out.update({kind + "_z0": z0, kind + "_zs": zs, kind + "_conf": conf, kind + "_means": np.array(means),
                    kind + "_covs": np.array(covs), kind + "_mp_mean": mm, kind + "_mp_cov": cc,
                    kind + "_meas": meas, **{kind + "_gate_" + k: v for k, v in gate.items()}})
    np.savez_compressed(os.path.join(GOLD, "ref_kf_cv8.npz"), **out)
    print("kf cv8 vectors", {k: v.shape for k, v in out.items() if k.startswith("xyah")})


# (case, stream id, K, frames, emb_dim, tracker kwargs); MOT17 cases run without ReID
BOTSORT_CASES = [
    ("synth_k40_e64", 10, 40, 80, 64, {}),
    ("synth_k100_e128", 11, 100, 50, 128, {}),
    ("synth_k30_e32_fuse", 12, 30, 80, 32, {"fuse_first_associate": True, "appearance_thresh": 0.4}),
    ("synth_k50_noreid", 13, 50, 80, 0, {"with_reid": False}),
    ("MOT17_02_FRCNN", None, 0, 0, 0, {"with_reid": False}),
]


# BoT-SORT with similarity camera warps (rotation): multi_gmc's kron(I4, R) couples x<->y, w<->h
BOTSORT_ROT_CASES = [
    ("synth_k30_e32_rot", 14, 30, 60, 32, {}),
    ("synth_k40_noreid_rot", 15, 40, 50, 0, {"with_reid": False, "track_buffer": 10}),
]


def main_botsort_rot():
    for name, sid, k, f, e, kw in BOTSORT_ROT_CASES:
        if e:
            fr, em = make_stream(sid, k, f, emb_dim=e)
        else:
            fr, em = make_stream(sid, k, f), None
        warps = boosttrack_warps(sid, f)
        warps[::6] = deepocsort_warps(sid, f)[::6]          # some rotation-free warps on dense tracks
        trk = ref_shim.make_tracker("botsort", **kw)
        out = []
        for i, d in enumerate(fr):
            trk.cmc.warps = [warps[i].copy()]
            r = np.asarray(trk.update(d.copy(), IMG, None if em is None else em[i].copy()), dtype=float)
            r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
            out.append(np.c_[np.full(len(r), i), r])
        out = np.vstack(out)
        np.savez_compressed(os.path.join(GOLD, "ref_botsort_%s.npz" % name), rows=out, kw=np.array(repr(kw)), warps=warps)
        print("botsort", name, out.shape, "ids", int(out[:, 5].max()))


def main_botsort():
    dets = {s: load_det_txt(s) for s in SEQS}
    for name, sid, k, f, e, kw in BOTSORT_CASES:
        if sid is None:
            fr, em = frames_of(dets[name.replace("_", "-")].astype(np.float32).astype(np.float64)), None
        elif e:
            fr, em = make_stream(sid, k, f, emb_dim=e)
        else:
            fr, em = make_stream(sid, k, f), None
        out = run_ref("botsort", fr, em, **kw)
        np.savez_compressed(os.path.join(GOLD, "ref_botsort_%s.npz" % name), rows=out, kw=np.array(repr(kw)))
        print("botsort", name, out.shape, "ids", int(out[:, 5].max()))


# (case, stream id, K, frames, emb_dim, warps, tracker kwargs)
DEEPOCSORT_CASES = [
    ("synth_k20_e32", 20, 20, 60, 32, False, {}),
    ("synth_k50_e64_awoff", 21, 50, 60, 64, False, {"aw_off": True}),
    ("synth_k40_e16_cmc", 22, 40, 60, 16, True, {}),
    ("synth_k60_e32_cmc_dt1", 23, 60, 50, 32, True, {"delta_t": 1, "max_age": 5, "det_thresh": 0.3}),
    ("synth_k30_noemb_giou", 24, 30, 50, 32, False, {"embedding_off": True, "asso_func": "giou"}),
    ("synth_k100_e128", 25, 100, 40, 128, False, {}),
    # similarity warps (rotation + scale + shift), what the tracker's own CMC (`sof`, deepocsort.py:311,
    # estimateAffinePartial2D) produces: the covariance's x / y blocks get coupled
    ("synth_k30_e32_rot", 27, 30, 60, 32, "sim", {}),
    ("synth_k50_e16_rot_age5", 28, 50, 50, 16, "sim", {"max_age": 5, "det_thresh": 0.3, "delta_t": 2}),
]


def deepocsort_warps(seed, n_frames):
    """Scale + shift camera warps (the CMC side input), one 2x3 matrix per frame."""
    rng = np.random.default_rng(seed)
    w = np.zeros((n_frames, 2, 3))
    w[:, 0, 0] = 1 + rng.normal(0, 0.01, n_frames)
    w[:, 1, 1] = 1 + rng.normal(0, 0.01, n_frames)
    w[:, :, 2] = rng.normal(0, 3, (n_frames, 2))
    return w


def main_deepocsort(only=None):
    for name, sid, k, f, e, use_warps, kw in DEEPOCSORT_CASES:
        if only and only not in name:
            continue
        fr, em = make_stream(sid, k, f, emb_dim=e)
        warps = (boosttrack_warps(sid, f) if use_warps == "sim" else deepocsort_warps(sid, f)) if use_warps else None
        if use_warps == "sim":
            warps[::7] = deepocsort_warps(sid, f)[::7]        # every 7th frame a rotation-free warp on dense tracks
        trk = ref_shim.make_tracker("deepocsort", **kw)
        out = []
        for i, d in enumerate(fr):
            if warps is not None:
                trk.cmc.warps = [warps[i].copy()]
            r = np.asarray(trk.update(d.copy(), IMG, em[i].copy()), dtype=float)
            r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
            out.append(np.c_[np.full(len(r), i), r])
        out = np.vstack(out)
        extra = {"warps": warps} if warps is not None else {}
        np.savez_compressed(os.path.join(GOLD, "ref_deepocsort_%s.npz" % name), rows=out, kw=np.array(repr(kw)),
                            **extra)
        print("deepocsort", name, out.shape, "ids", int(out[:, 5].max()))


# (case, stream id, K, frames, emb_dim, warps, crowd, tracker kwargs)
BOOSTTRACK_CASES = [
    ("synth_k20_e32", 40, 20, 60, 32, False, False, {}),
    ("synth_k50_e64_cmc", 41, 50, 60, 64, True, False, {}),
    ("synth_k40_plain", 42, 40, 60, 16, False, False,
     {"with_reid": False, "use_rich_s": False, "use_sb": False, "use_vt": False}),
    ("synth_k60_crowd_v2", 43, 60, 40, 32, False, True, {"s_sim_corr": True, "max_age": 10}),
    ("synth_k30_noduo", 44, 30, 50, 8, False, False, {"use_ecc": False, "use_duo_boost": False, "use_vt": False}),
    ("synth_k30_nodlo", 45, 30, 50, 8, False, False, {"use_dlo_boost": False, "min_hits": 1, "det_thresh": 0.4}),
]


def boosttrack_warps(seed, n_frames):
    """Similarity camera warps (rotation + scale + shift), one 2x3 matrix per frame."""
    rng = np.random.default_rng(seed)
    th, sc = rng.normal(0, 0.003, n_frames), 1 + rng.normal(0, 0.005, n_frames)
    w = np.zeros((n_frames, 2, 3))
    w[:, 0, 0] = sc * np.cos(th); w[:, 0, 1] = -sc * np.sin(th)
    w[:, 1, 0] = sc * np.sin(th); w[:, 1, 1] = sc * np.cos(th)
    w[:, :, 2] = rng.normal(0, 3, (n_frames, 2))
    return w


def main_boosttrack():
    for name, sid, k, f, e, use_warps, crowd, kw in BOOSTTRACK_CASES:
        fr, em = make_stream(sid, k, f, emb_dim=e, crowd=crowd)
        warps = boosttrack_warps(sid, f) if use_warps else None
        trk = ref_shim.make_tracker("boosttrack", **kw)
        out = []
        for i, d in enumerate(fr):
            if warps is not None:
                trk.cmc.warps = [warps[i].copy()]
            r = np.asarray(trk.update(d.copy(), IMG, em[i].copy()), dtype=float)
            r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
            out.append(np.c_[np.full(len(r), i), r])
        out = np.vstack(out)
        extra = {"warps": warps} if warps is not None else {}
        np.savez_compressed(os.path.join(GOLD, "ref_boosttrack_%s.npz" % name), rows=out, kw=np.array(repr(kw)),
                            **extra)
        print("boosttrack", name, out.shape, "ids", int(out[:, 5].max()))


def main_mot17_emb(seqs=("MOT17-02-FRCNN",)):
    """DeepOCSORT / BoostTrack (and, for MOT17-04, BoT-SORT) on MOT17-mini det.txt with seeded embeddings."""
    for seq in seqs:
        dets = load_det_txt(seq).astype(np.float32).astype(np.float64)
        fr = frames_of(dets)
        em = mot17_embs(fr, 48, 77)
        name = seq.replace("-", "_")
        todo = [("deepocsort", {"det_thresh": 0.3}), ("boosttrack", {"det_thresh": 0.5})]
        if seq != "MOT17-02-FRCNN":
            todo.append(("botsort", {}))                       # with ReID on the seeded embeddings
        for tracker, kw in todo:
            out = run_ref(tracker, fr, em, **kw)
            np.savez_compressed(os.path.join(GOLD, "ref_%s_%s.npz" % (tracker, name)), rows=out, kw=np.array(repr(kw)))
            print(tracker, name, out.shape, "ids", int(out[:, 5].max()))


def strongsort_gate_vectors(la, rng):
    """Known-answer vectors of gated_metric's post-processing (gate_cost_matrix + id-preservation
    weighting, row a22), _enhance_cost_matrix and iou_cost, from the reference's own functions on
    the reference's own Track / Detection objects (tracks driven through predict / update so that
    covariances, histories and scores are the ones the state machine produces; a few attributes are
    then overwritten to reach every branch)."""
    import types
    from oracle import strongsort as so
    cmc = sys.modules.setdefault("boxmot.motion.cmc", types.ModuleType("boxmot.motion.cmc"))
    if not hasattr(cmc, "get_cmc_method"):
        cmc.get_cmc_method = lambda _name: (lambda: None)
    from boxmot.trackers.strongsort.sort import iou_matching
    from boxmot.trackers.strongsort.sort.detection import Detection
    from boxmot.trackers.strongsort.sort.track import Track
    from boxmot.trackers.strongsort.sort.tracker import Tracker
    T, D, E = 40, 57, 16

    def rand_det(c=None, cls=None):
        wh = rng.uniform(20, 200, 2)
        c = rng.random(2) * [1800, 1000] if c is None else c
        d = Detection(np.r_[c - wh / 2, wh], float(rng.uniform(0.05, 0.99)),
                      float(rng.integers(0, 3)) if cls is None else cls, 0, rng.normal(0, 1, E).astype(np.float32))
        d.quality_score = float(rng.uniform(0.1, 1.1))
        return d

    tracks = []
    for t in range(T):
        d = rand_det()
        trk = Track(d, t + 1, 2, 50, 0.9)
        for _ in range(int(rng.integers(0, 14))):
            trk.predict()
            if rng.random() < 0.8:
                c = trk.mean[:2] + rng.normal(0, 4, 2)
                trk.update(rand_det(c, trk.cls))
        if t % 5 != 4:
            trk.predict()          # the state gated_metric sees: predicted, time_since_update >= 1
        if t % 7 == 0:
            trk.appearance_consistency = float(rng.uniform(0.1, 0.5))
        if t % 6 == 0:
            trk.low_conf_streak = 3
        if t % 9 == 0:
            trk.quality_score = 0.25
        tracks.append(trk)
    def near(k):       # a detection the gate accepts for track k (most of the time): its own box, jittered
        h = k.mean[3] * (1 + rng.normal(0, 0.03))
        w = k.mean[2] * k.mean[3] * (1 + rng.normal(0, 0.03))
        d = rand_det(k.mean[:2] + rng.normal(0, 0.04, 2) * k.mean[3])
        d.tlwh = np.r_[d.tlwh[:2] + d.tlwh[2:] / 2 - [w / 2, h / 2], w, h]
        return d

    dets = [near(tracks[i % T]) if i < 45 else rand_det() for i in range(D)]
    tidx, didx = list(range(T)), list(range(D))
    cost0 = rng.uniform(0, 0.4, (T, D))
    cost0[3] = 1e5
    g = la.gate_cost_matrix(cost0.copy(), tracks, dets, tidx, didx, 0.995)
    g = Tracker._apply_id_preservation_weighting(types.SimpleNamespace(id_preservation_weight=0.1), g, tracks, dets,
                                                 tidx, didx)
    enh = la._enhance_cost_matrix(cost0.copy(), tracks, dets, tidx, didx)
    iou = iou_matching.iou_cost(tracks, dets, tidx, didx)
    trk = np.zeros((T, so.TRK_FIELDS))
    for t, k in enumerate(tracks):
        v = k.velocity_history[-1] if k.velocity_history else np.zeros(2)
        trk[t] = [k.age, k.hits, k.time_since_update, k.motion_consistency, k.appearance_consistency,
                  k.quality_score, k.high_conf_streak, k.low_conf_streak, float(len(k.velocity_history) > 0),
                  v[0], v[1], k.conf, k.cls]
    print("strongsort gate: gated", int((g > 500).sum()), "of", g.size, "iou<1", int((iou < 1).sum()))
    return dict(g_means=np.array([k.mean for k in tracks]), g_covs=np.array([k.covariance for k in tracks]),
                g_trk=trk, g_tlwh=np.array([d.tlwh for d in dets]), g_dq=np.array([d.quality_score for d in dets]),
                g_dconf=np.array([d.conf for d in dets]), g_dcls=np.array([d.cls for d in dets]), g_cost0=cost0,
                g_gated=g, g_enh=enh, g_iou=iou, g_lambda=np.array(0.995), g_idw=np.array(0.1))


def strongsort_feature_vectors(rng):
    """Known-answer vectors of the feature EMA inside the reference's Track.update (track.py:238-268):
    reference Track objects updated with detections whose features are a noisy copy of the track's
    identity (similar), an unrelated vector (dissimilar) or something in between, at low / medium /
    high confidence -- every branch of _calculate_adaptive_ema_alpha."""
    from boxmot.trackers.strongsort.sort.detection import Detection
    from boxmot.trackers.strongsort.sort.track import Track
    N, E, steps = 48, 64, 5
    ident = rng.normal(0, 1, (N, E))
    last, feat, conf, ac0, out, ac1 = [], [], [], [], [], []
    tracks = []
    for n in range(N):
        d = Detection(np.array([100.0, 100.0, 50.0, 120.0]), 0.9, 0.0, 0, (ident[n] * 3.0).astype(np.float32))
        tracks.append(Track(d, n + 1, 2, 50, 0.9))
    for k in range(steps):
        for n, trk in enumerate(tracks):
            mix = [0.05, 0.8, 3.0][(n + k) % 3]
            f = (ident[n] + mix * rng.normal(0, 1, E)).astype(np.float32) * float(rng.uniform(0.5, 4))
            c = float([0.95, 0.5, 0.15][(n // 3 + k) % 3])
            d = Detection(np.array([100.0, 100.0, 50.0, 120.0]), c, 0.0, 0, f.copy())
            last.append(np.asarray(trk.features[-1], dtype=np.float64))
            feat.append(f)
            conf.append(c)
            ac0.append(trk.appearance_consistency)
            trk.predict()
            trk.update(d)
            out.append(np.asarray(trk.features[-1], dtype=np.float64))
            ac1.append(trk.appearance_consistency)
    print("strongsort feature EMA:", len(out), "updates, consistency range", min(ac1), max(ac1))
    return dict(f_last=np.array(last), f_feat=np.array(feat), f_conf=np.array(conf), f_ac0=np.array(ac0),
                f_out=np.array(out), f_ac1=np.array(ac1), f_alpha=np.array(0.9))


def strongsort_track_row(k):
    """A reference Track as a row of oracle.strongsort's TS_* table."""
    from oracle import strongsort as so
    a = np.zeros(so.TS_FIELDS)
    v = k.velocity_history[-1] if k.velocity_history else np.zeros(2)
    a[:13] = [k.age, k.hits, k.time_since_update, k.motion_consistency, k.appearance_consistency, k.quality_score,
              k.high_conf_streak, k.low_conf_streak, len(k.velocity_history), v[0], v[1], k.conf, k.cls]
    a[so.TS_STATE], a[so.TS_STABILITY] = k.state, k.stability_score
    a[so.TS_CONFIRMED], a[so.TS_MISSED] = k.confirmed_detections, k.missed_detections
    a[so.TS_NPOS] = len(k.position_history)
    a[so.TS_PX:so.TS_PY + 1] = k.position_history[-1]
    a[so.TS_NCONF] = len(k.confidence_history)
    a[so.TS_CONF0:so.TS_CONF0 + len(k.confidence_history)] = k.confidence_history
    a[so.TS_DETIND] = k.det_ind
    return a


def strongsort_track_vectors(rng):
    """Known-answer vectors of the reference's Track.predict / update / mark_missed (track.py:195-317):
    64 Track objects stepped for 30 frames (update with p = 0.75, else a miss), different n_init /
    max_age, confidences from all three bands; every call is recorded as (op, state before, detection,
    state after).  Detections carry no feature (the feature step is its own unit); the appearance
    consistency that step would have produced is set by hand before each update."""
    from boxmot.trackers.strongsort.sort.detection import Detection
    from boxmot.trackers.strongsort.sort.track import Track
    rec = {k: [] for k in ("op", "m0", "c0", "a0", "det", "par", "m1", "c1", "a1")}
    for n in range(64):
        n_init, max_age = [2, 3, 5][n % 3], [3, 6, 50][n % 3 if n % 2 else (n + 1) % 3]
        c = rng.random(2) * [1800, 1000]
        wh = rng.uniform(20, 200, 2)
        d = Detection(np.r_[c - wh / 2, wh], float(rng.uniform(0.05, 0.99)), float(n % 3), n, None)
        d.quality_score = float(rng.uniform(0.1, 1.0))
        trk = Track(d, n + 1, n_init, max_age, 0.9)
        vel = rng.normal(0, 4, 2)
        for f in range(30):
            ops = ["camera", "predict", "update" if rng.random() < 0.75 else "miss"]
            for op in ops:
                m0, c0, a0 = trk.mean.copy(), trk.covariance.copy(), strongsort_track_row(trk)
                det = np.zeros(7)
                if op == "predict":
                    trk.predict()
                elif op == "camera":
                    th, sc = rng.normal(0, 0.01), 1 + rng.normal(0, 0.01)
                    wm = np.array([[sc * np.cos(th), -sc * np.sin(th), rng.normal(0, 4)],
                                   [sc * np.sin(th), sc * np.cos(th), rng.normal(0, 4)]])
                    det = np.r_[wm.ravel(), 0.0]
                    trk.camera_update(wm)
                elif op == "update":
                    c = c + vel + rng.normal(0, 1.5, 2)
                    wh = wh * (1 + rng.normal(0, 0.02, 2))
                    band = rng.integers(0, 3)
                    conf = float(rng.uniform(*[(0.05, 0.29), (0.31, 0.69), (0.71, 0.99)][band]))
                    dd = Detection(np.r_[c - wh / 2, wh], conf, float(rng.integers(0, 3)), int(rng.integers(0, 50)), None)
                    trk.appearance_consistency = float(rng.uniform(0.2, 1.0))
                    a0 = strongsort_track_row(trk)
                    det = np.r_[dd.to_xyah(), conf, dd.cls, dd.det_ind]
                    trk.update(dd)
                else:
                    trk.mark_missed()
                for k, v in zip(("op", "m0", "c0", "a0", "det", "par", "m1", "c1", "a1"),
                                ({"predict": 0, "update": 1, "miss": 2, "camera": 3}[op], m0, c0, a0, det, [n_init, max_age],
                                 trk.mean.copy(), trk.covariance.copy(), strongsort_track_row(trk))):
                    rec[k].append(v)
            if trk.is_deleted():
                break
    out = {"t_" + k: np.array(v) for k, v in rec.items()}
    st = out["t_a1"][:, 13]
    print("strongsort track ops:", np.bincount(out["t_op"]), "final states", np.bincount(st.astype(int)))
    return out


def strongsort_cascade_vectors(la, rng):
    """The reference's matching_cascade (linear_assignment.py:90-171) with the gated_metric of
    tracker.py:186-206 on 40 reference Track objects at time_since_update 1..4 with a filled sample
    bank, against 60 detections (most of them a jittered copy of a track's box and identity)."""
    import types
    from boxmot.trackers.strongsort.sort.detection import Detection
    from boxmot.trackers.strongsort.sort.track import Track
    from boxmot.trackers.strongsort.sort.tracker import Tracker
    T, D, E, CAP = 40, 60, 32, 12
    ident = rng.normal(0, 1, (T, E))
    metric = la.NearestNeighborDistanceMetric("cosine", 0.3, None)
    tracks = []
    for t in range(T):
        wh = rng.uniform(30, 150, 2) * [1, 2]
        c = rng.random(2) * [1800, 1000]
        d = Detection(np.r_[c - wh / 2, wh], float(rng.uniform(0.3, 0.99)), float(t % 2), t, None)
        d.quality_score = float(rng.uniform(0.2, 1.0))
        trk = Track(d, t + 1, 2, 50, 0.9)
        vel = rng.normal(0, 3, 2)
        for _ in range(int(rng.integers(2, 9))):
            trk.predict()
            c = c + vel + rng.normal(0, 1, 2)
            dd = Detection(np.r_[c - wh / 2, wh], float(rng.uniform(0.2, 0.99)), float(t % 2), 0, None)
            trk.appearance_consistency = float(rng.uniform(0.3, 1.0))
            trk.update(dd)
        for _ in range(1 + t % 4):
            trk.predict()                       # time_since_update 1..4: four cascade levels
        nb = int(rng.integers(0, CAP + 1)) if t % 9 else 0
        metric.samples[trk.id] = [((ident[t] + 0.1 * rng.normal(0, 1, E)) * rng.uniform(0.5, 2)).astype(np.float32)
                                  for _ in range(nb)]
        tracks.append(trk)
    dets = []
    for i in range(D):
        if i < 48:
            k = tracks[i % T]
            h = k.mean[3] * (1 + rng.normal(0, 0.02))
            w = k.mean[2] * k.mean[3] * (1 + rng.normal(0, 0.02))
            c = k.mean[:2] + rng.normal(0, 0.03, 2) * k.mean[3]
            f = ident[i % T] + (0.1 if i % 5 else 1.5) * rng.normal(0, 1, E)
        else:
            w, h = rng.uniform(30, 150), rng.uniform(60, 300)
            c = rng.random(2) * [1800, 1000]
            f = rng.normal(0, 1, E)
        d = Detection(np.r_[c[0] - w / 2, c[1] - h / 2, w, h], float(rng.uniform(0.05, 0.99)), float(i % 2), i,
                      (f * rng.uniform(0.5, 2)).astype(np.float32))
        d.quality_score = float(rng.uniform(0.1, 1.1))
        dets.append(d)
    fake = types.SimpleNamespace(id_preservation_weight=0.1)

    def gated_metric(trs, des, ti, di):
        cost = metric.distance(np.array([des[i].feat for i in di]), np.array([trs[i].id for i in ti]))
        cost = la.gate_cost_matrix(cost, trs, des, ti, di, 0.995)
        return Tracker._apply_id_preservation_weighting(fake, cost, trs, des, ti, di)

    tidx, didx = list(range(T)), list(range(D))
    m, ut, ud = la.matching_cascade(gated_metric, 0.3, 3, tracks, dets, tidx, didx)
    bank = np.zeros((T, CAP, E), np.float32)
    nbs = np.zeros(T, np.int32)
    for t, k in enumerate(tracks):
        sm = metric.samples[k.id]
        nbs[t] = len(sm)
        if sm:
            bank[t, : len(sm)] = np.array(sm)
    tlwh = np.array([d.tlwh for d in dets])
    print("strongsort cascade:", len(m), "matches,", len(ut), "unmatched tracks,", len(ud), "unmatched detections")
    return dict(c_means=np.array([k.mean for k in tracks]), c_covs=np.array([k.covariance for k in tracks]),
                c_trk=np.array([strongsort_track_row(k) for k in tracks]), c_bank=bank, c_nb=nbs, c_tlwh=tlwh,
                c_dattr=np.array([[d.quality_score, d.conf, d.cls] for d in dets]),
                c_feats=np.array([d.feat for d in dets]), c_matches=np.array(sorted(m)).reshape(-1, 2),
                c_ut=np.array(sorted(ut)), c_ud=np.array(sorted(ud)), c_maxd=np.array(0.3), c_depth=np.array(3))


def main_strongsort_units():
    """Known-answer vectors of StrongSORT's appearance metric and min_cost_matching, produced by the
    reference's own functions (rows a18 / a21)."""
    ref_shim.load()
    from boxmot.trackers.strongsort.sort import linear_assignment as la
    rng = np.random.default_rng(21)
    T, D, E, B = 12, 37, 96, 20
    metric = la.NearestNeighborDistanceMetric("cosine", 0.2, None)
    nb = rng.integers(0, B + 1, T)
    nb[3] = 0
    bank = (rng.normal(0, 1, (T, B, E)) * rng.uniform(0.5, 2, (T, B, 1))).astype(np.float32)
    feats = rng.normal(0, 1, (D, E)).astype(np.float32)
    # a few near-duplicates so that the minimum is small somewhere
    for t in range(T):
        if nb[t]:
            feats[t] = bank[t, rng.integers(0, nb[t])] * 1.7 + rng.normal(0, 0.05, E)
    for t in range(T):
        if nb[t]:
            metric.samples[t] = [bank[t, i] for i in range(nb[t])]
    nn = metric.distance(feats, list(range(T)))

    class _T:       # neutral attributes: _enhance_cost_matrix multiplies by exactly 1.0
        quality_score = 0.5
        conf = 0.5

    class _D:
        quality_score = 0.5
        conf = 0.5

    P, N = 48, 30
    costs = np.zeros((P, N, N))
    nrs, ncs = rng.integers(1, N + 1, P), rng.integers(1, N + 1, P)
    xs = -np.ones((P, N), dtype=np.int64)
    for p in range(P):
        c = rng.random((nrs[p], ncs[p]))
        costs[p, : nrs[p], : ncs[p]] = c
        m, _, _ = la.min_cost_matching(lambda tr, de, ti, di, c=c: c.copy(), 0.5, [_T()] * nrs[p], [_D()] * ncs[p])
        for r, cc in m:
            xs[p, r] = cc
    gate = strongsort_gate_vectors(la, rng)
    gate.update(strongsort_feature_vectors(rng))
    gate.update(strongsort_track_vectors(rng))
    gate.update(strongsort_cascade_vectors(la, rng))
    np.savez_compressed(os.path.join(GOLD, "ref_strongsort_units.npz"), bank=bank, nb=nb, feats=feats, nn=nn,
                        costs=costs, nrs=nrs, ncs=ncs, xs=xs, max_distance=np.array(0.5), **gate)
    print("strongsort units: nn", nn.shape, "min", nn.min(), "matching problems", P, "pairs", int((xs >= 0).sum()))


# (case, stream id, K, frames, emb_dim, warps, tracker kwargs)
STRONGSORT_CASES = [
    ("synth_k20_e32", 60, 20, 60, 32, False, {}),
    ("synth_k30_e64_b20", 61, 30, 70, 64, False, {"nn_budget": 20}),
    ("synth_k40_e32_cmc", 62, 40, 50, 32, True, {"nn_budget": 50}),
    ("synth_k25_e96_age8", 63, 25, 80, 96, False, {"max_age": 8, "nn_budget": 12}),
    ("synth_k60_e32_crowd", 64, 60, 40, 32, False, {"nn_budget": 30, "max_age": 10}),
    ("synth_k15_e32_ninit3", 65, 15, 90, 32, True, {"n_init": 3, "max_age": 6, "nn_budget": 8, "min_conf": 0.3}),
    ("synth_k50_e128_hi", 66, 50, 40, 128, False, {"conf_thresh_high": 0.5, "conf_thresh_low": 0.2, "max_cos_dist": 0.3}),
]


def strongsort_inputs(sid, k, f, e, crowd=False):
    return strongsort_stream(sid, k, f, e, crowd)


def main_strongsort():
    """StrongSORT end to end in its runnable configuration (ref_shim._make_strongsort: R4, M1, M2, R5)."""
    for name, sid, k, f, e, use_warps, kw in STRONGSORT_CASES:
        fr, em = strongsort_inputs(sid, k, f, e, crowd="crowd" in name)
        warps = boosttrack_warps(sid, f) if use_warps else None
        trk = ref_shim.make_tracker("strongsort", **kw)
        out, ntr = [], []
        for i, d in enumerate(fr):
            trk.cmc.warps = [warps[i].copy()] if warps is not None else []
            r = np.asarray(trk.update(d.copy(), IMG, em[i].copy()), dtype=float)
            r = r.reshape(-1, r.shape[-1])[:, :8] if r.size else np.empty((0, 8))
            out.append(np.c_[np.full(len(r), i), r])
            ntr.append((len(trk.tracker.tracks), len(trk.tracker.lost_tracks), trk.tracker._next_id))
        out = np.vstack(out)
        extra = {"warps": warps} if warps is not None else {}
        np.savez_compressed(os.path.join(GOLD, "ref_strongsort_%s.npz" % name), rows=out, kw=np.array(repr(kw)),
                            ntracks=np.array(ntr), **extra)
        print("strongsort", name, out.shape, "ids", int(out[:, 5].max()), "tracks/lost/next_id at end", ntr[-1],
              "recovered", sum(1 for a, b in zip(ntr, ntr[1:]) if b[1] < a[1]))


def main_mot_writer():
    """The reference's own convert_to_mot_format + write_mot_results (boxmot/engine/utils.py:101-172) on
    the ByteTrack MOT17-02-FRCNN golden rows -> tests/golden/ref_mot_results_bytetrack_MOT17_02_FRCNN.txt.
    engine/utils.py imports ultralytics (absent), so the two functions are executed from their source text;
    R6: the fork's rewritten utils/ops.py has no `xyxy2ltwh` (convert_to_mot_format raises AttributeError as
    shipped); it is the same map as the fork's own `xyxy2tlwh` (x1, y1, x2 - x1, y2 - y1), injected here."""
    import tempfile
    from pathlib import Path
    ref_shim.load()
    from boxmot.utils import ops
    if not hasattr(ops, "xyxy2ltwh"):
        ops.xyxy2ltwh = ops.xyxy2tlwh                      # R6
    src = open(os.path.join(ref_shim.REF_ROOT, "boxmot", "engine", "utils.py")).read()
    src = src[src.index("def convert_to_mot_format("):].split("\ndef eval_setup")[0]
    g = {"ops": ops, "np": np}
    exec("from pathlib import Path\nfrom typing import Union\nimport torch\nResults = object\n" + src, g)
    z = np.load(os.path.join(GOLD, "ref_bytetrack_MOT17_02_FRCNN.npz"))["rows"]
    path = Path(tempfile.mkdtemp()) / "MOT17-02-FRCNN.txt"
    for f in range(int(z[:, 0].max()) + 1):
        r = z[z[:, 0] == f][:, 1:]
        if r.size:
            g["write_mot_results"](path, g["convert_to_mot_format"](r, f + 1))
    txt = path.read_text()
    open(os.path.join(GOLD, "ref_mot_results_bytetrack_MOT17_02_FRCNN.txt"), "w").write(txt)
    print("mot writer:", len(txt.splitlines()), "rows;", txt.splitlines()[0])


def main():
    os.makedirs(GOLD, exist_ok=True)
    dets = {s: load_det_txt(s) for s in SEQS}
    np.savez_compressed(os.path.join(GOLD, "mot17_mini_dets.npz"),
                        **{s.replace("-", "_"): dets[s].astype(np.float32) for s in SEQS})
    for tracker in ("ocsort", "bytetrack"):
        for s in SEQS:
            # float32 storage is lossless: det.txt has <= 6 significant digits and the
            # reference rounds inputs to float32 anyway (basetracker.py:122-128)
            fr = frames_of(dets[s].astype(np.float32).astype(np.float64))
            out = run_ref(tracker, fr)
            np.savez_compressed(os.path.join(GOLD, "ref_%s_%s.npz" % (tracker, s.replace("-", "_"))), rows=out)
            print(tracker, s, out.shape)
        for name, mk, kw in synth_cases():
            if tracker == "bytetrack":
                kw = {}
            out = run_ref(tracker, mk(), **kw)
            np.savez_compressed(os.path.join(GOLD, "ref_%s_%s.npz" % (tracker, name)), rows=out,
                                kw=np.array(repr(kw)))
            print(tracker, name, out.shape)


if __name__ == "__main__":
    if "--per-class" in sys.argv:
        main_per_class()
    elif "--deepocsort-per-class" in sys.argv:
        main_deepocsort_per_class()
    elif "--per-class-shared" in sys.argv:
        main_per_class_shared()
    elif "--kf" in sys.argv:
        main_kf()
    elif "--botsort-rot" in sys.argv:
        main_botsort_rot()
    elif "--botsort" in sys.argv:
        main_botsort()
    elif "--deepocsort-rot" in sys.argv:
        main_deepocsort(only="_rot")
    elif "--deepocsort" in sys.argv:
        main_deepocsort()
    elif "--boosttrack" in sys.argv:
        main_boosttrack()
    elif "--mot17-04-emb" in sys.argv:
        main_mot17_emb(("MOT17-04-FRCNN",))
    elif "--mot17-emb" in sys.argv:
        main_mot17_emb()
    elif "--strongsort" in sys.argv:
        main_strongsort()
    elif "--mot-writer" in sys.argv:
        main_mot_writer()
    elif "--strongsort-units" in sys.argv:
        main_strongsort_units()
    else:
        main()
