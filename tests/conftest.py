import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def mot17_frames(seq):
    """Per-frame (N,6) detection arrays of a MOT17-mini sequence (fixture copy of
    assets/MOT17-mini/train/<seq>/det/det.txt, see oracle/gen_golden.py)."""
    rows = np.load(os.path.join(GOLD, "mot17_mini_dets.npz"))[seq].astype(np.float64)
    n = int(rows[:, 0].max())
    return [np.c_[rows[rows[:, 0] == f][:, 1:], np.zeros(int((rows[:, 0] == f).sum()))]
            for f in range(1, n + 1)]


def golden_rows(tracker, case):
    z = np.load(os.path.join(GOLD, "ref_%s_%s.npz" % (tracker, case)))
    rows = z["rows"]
    kw = eval(str(z["kw"])) if "kw" in z.files else {}
    return rows, kw


def split_by_frame(rows, n_frames):
    return [rows[rows[:, 0] == f][:, 1:] for f in range(n_frames)]


SYNTH_CASES = {
    "synth_k100": (0, 100, 80),
    "synth_k30_byte": (1, 30, 120),
    "synth_k30_giou": (2, 30, 80),
    "synth_k50_dt1": (3, 50, 80),
    "synth_k40_ciou": (4, 40, 60),
}
OCSORT_YAML = dict(min_conf=0.1, det_thresh=0.6, max_age=30, min_hits=3, asso_threshold=0.3,
                   delta_t=3, asso_func="iou", inertia=0.1, use_byte=False, Q_xy_scaling=0.01,
                   Q_s_scaling=0.0001)


BYTETRACK_YAML = dict(min_conf=0.1, track_thresh=0.6, match_thresh=0.9, track_buffer=30, frame_rate=30)


BOTSORT_YAML = dict(track_high_thresh=0.6, track_low_thresh=0.1, new_track_thresh=0.7, track_buffer=30,
                    match_thresh=0.8, proximity_thresh=0.5, appearance_thresh=0.25, frame_rate=30,
                    fuse_first_associate=False, with_reid=True)
# case -> (stream id, K, frames, emb_dim); see oracle/gen_golden.py BOTSORT_CASES
BOTSORT_CASES = {
    "synth_k40_e64": (10, 40, 80, 64),
    "synth_k100_e128": (11, 100, 50, 128),
    "synth_k30_e32_fuse": (12, 30, 80, 32),
    "synth_k50_noreid": (13, 50, 80, 0),
    "synth_k30_e32_rot": (14, 30, 60, 32),           # similarity camera warps: multi_gmc with rotation (dense covariance)
    "synth_k40_noreid_rot": (15, 40, 50, 0),
    "MOT17_02_FRCNN": None,
    "MOT17_04_FRCNN": None,      # real det.txt rows + seeded 48-d embeddings, with ReID (gen_golden --mot17-04-emb)
}


def botsort_warps(case):
    z = np.load(os.path.join(GOLD, "ref_botsort_%s.npz" % case))
    return z["warps"] if "warps" in z.files else None


def botsort_case(case):
    """(frames, embs or None, tracker kwargs, golden rows) of a BoT-SORT fixture."""
    from boxmot_b200.synth import make_stream
    rows, kw = golden_rows("botsort", case)
    spec = BOTSORT_CASES[case]
    if case == "MOT17_04_FRCNN":
        fr, em = _mot17_emb_case(case)
        return fr, em, kw, rows
    if spec is None:
        return mot17_frames(case), None, kw, rows
    sid, k, f, e = spec
    if e:
        fr, em = make_stream(sid, k, f, emb_dim=e)
        return fr, em, kw, rows
    return make_stream(sid, k, f), None, kw, rows


DEEPOCSORT_YAML = dict(det_thresh=0.5, max_age=30, min_hits=3, iou_threshold=0.3, delta_t=3, asso_func="iou",
                       inertia=0.2, w_association_emb=0.75, alpha_fixed_emb=0.95, aw_param=0.5,
                       embedding_off=False, cmc_off=False, aw_off=False, Q_xy_scaling=0.01, Q_s_scaling=0.0001)
# case -> (stream id, K, frames, emb_dim); see oracle/gen_golden.py DEEPOCSORT_CASES
DEEPOCSORT_CASES = {
    "synth_k20_e32": (20, 20, 60, 32),
    "synth_k50_e64_awoff": (21, 50, 60, 64),
    "synth_k40_e16_cmc": (22, 40, 60, 16),
    "synth_k60_e32_cmc_dt1": (23, 60, 50, 32),
    "synth_k30_noemb_giou": (24, 30, 50, 32),
    "synth_k100_e128": (25, 100, 40, 128),
    "synth_k30_e32_rot": (27, 30, 60, 32),           # similarity warps: rotation couples x and y (dense covariance)
    "synth_k50_e16_rot_age5": (28, 50, 50, 16),
    "MOT17_02_FRCNN": None,      # real det.txt rows + seeded 48-d embeddings (gen_golden --mot17-emb)
    "MOT17_04_FRCNN": None,
}


def _mot17_emb_case(case):
    from boxmot_b200.synth import mot17_embs
    fr = [d.astype(np.float32).astype(np.float64) for d in mot17_frames(case)]
    return fr, mot17_embs(fr, 48, 77)


def deepocsort_case(case):
    """(frames, embs, warps or None, tracker kwargs, golden rows) of a DeepOCSORT fixture."""
    from boxmot_b200.synth import make_stream
    z = np.load(os.path.join(GOLD, "ref_deepocsort_%s.npz" % case))
    if case.startswith("MOT17"):
        fr, em = _mot17_emb_case(case)
    else:
        sid, k, f, e = DEEPOCSORT_CASES[case]
        fr, em = make_stream(sid, k, f, emb_dim=e)
    return fr, em, (z["warps"] if "warps" in z.files else None), eval(str(z["kw"])), z["rows"]


def deepocsort_per_class_case():
    """(frames, embs, warps, golden rows) of the per_class DeepOCSORT fixture: three classes by image
    column (oracle/gen_golden.py per_class_emb_frames)."""
    from boxmot_b200.synth import make_stream
    z = np.load(os.path.join(GOLD, "ref_deepocsort_synth_k40_perclass.npz"))
    fr, em = make_stream(26, 40, 60, emb_dim=32)
    out = []
    for d in fr:
        d = d.copy()
        d[:, 5] = np.floor((d[:, 0] + d[:, 2]) / 2.0 / 640.0).clip(0, 2)
        out.append(d)
    return out, em, z["warps"], z["rows"]


BOOSTTRACK_YAML = dict(max_age=60, min_hits=3, det_thresh=0.6, iou_threshold=0.3, use_ecc=True, min_box_area=10,
                       aspect_ratio_thresh=1.6, lambda_iou=0.5, lambda_mhd=0.25, lambda_shape=0.25,
                       use_dlo_boost=True, use_duo_boost=True, dlo_boost_coef=0.65, s_sim_corr=False,
                       use_rich_s=True, use_sb=True, use_vt=True, with_reid=True)
# case -> (stream id, K, frames, emb_dim, crowd); see oracle/gen_golden.py BOOSTTRACK_CASES
BOOSTTRACK_CASES = {
    "synth_k20_e32": (40, 20, 60, 32, False),
    "synth_k50_e64_cmc": (41, 50, 60, 64, False),
    "synth_k40_plain": (42, 40, 60, 16, False),
    "synth_k60_crowd_v2": (43, 60, 40, 32, True),
    "synth_k30_noduo": (44, 30, 50, 8, False),
    "synth_k30_nodlo": (45, 30, 50, 8, False),
    "MOT17_02_FRCNN": None,      # real det.txt rows + seeded 48-d embeddings (gen_golden --mot17-emb)
    "MOT17_04_FRCNN": None,
}


def boosttrack_case(case):
    """(frames, embs, warps or None, tracker kwargs, golden rows) of a BoostTrack fixture."""
    from boxmot_b200.synth import make_stream
    z = np.load(os.path.join(GOLD, "ref_boosttrack_%s.npz" % case))
    if case.startswith("MOT17"):
        fr, em = _mot17_emb_case(case)
    else:
        sid, k, f, e, crowd = BOOSTTRACK_CASES[case]
        fr, em = make_stream(sid, k, f, emb_dim=e, crowd=crowd)
    return fr, em, (z["warps"] if "warps" in z.files else None), eval(str(z["kw"])), z["rows"]


# the constructor defaults of the reference's StrongSort that strongsort.yaml also carries
STRONGSORT_YAML = dict(min_conf=0.1, ema_alpha=0.9, max_cos_dist=0.15, max_iou_dist=0.7, max_age=50, n_init=2,
                       mc_lambda=0.995, nn_budget=150, conf_thresh_high=0.7, conf_thresh_low=0.3,
                       id_preservation_weight=0.1)
# case -> (stream id, K, frames, emb_dim); see oracle/gen_golden.py STRONGSORT_CASES
STRONGSORT_CASES = {
    "synth_k20_e32": (60, 20, 60, 32),
    "synth_k30_e64_b20": (61, 30, 70, 64),
    "synth_k40_e32_cmc": (62, 40, 50, 32),
    "synth_k25_e96_age8": (63, 25, 80, 96),
    "synth_k60_e32_crowd": (64, 60, 40, 32),
    "synth_k15_e32_ninit3": (65, 15, 90, 32),
    "synth_k50_e128_hi": (66, 50, 40, 128),
}


def strongsort_case(case):
    """(frames, embs, warps or None, tracker kwargs, golden rows) of a StrongSORT fixture."""
    from boxmot_b200.synth import strongsort_stream
    z = np.load(os.path.join(GOLD, "ref_strongsort_%s.npz" % case))
    sid, k, f, e = STRONGSORT_CASES[case]
    fr, em = strongsort_stream(sid, k, f, e, crowd="crowd" in case)
    return fr, em, (z["warps"] if "warps" in z.files else None), eval(str(z["kw"])), z["rows"]


def case_frames(case):
    from boxmot_b200.synth import make_stream
    if case.startswith("MOT17"):
        return mot17_frames(case)
    sid, k, f = SYNTH_CASES[case]
    return make_stream(sid, k, f)


def assert_rows_match(got, want, frame, box_rtol=1e-6, conf_rtol=0.0):
    """ids / conf / cls / det_ind exact and in the reference's row order; boxes 1e-6 rel.
    conf_rtol > 0: the reported confidence is a COMPUTED value (BoostTrack's boosts go through
    exp / pow, whose last bit differs between libm, numpy's SIMD kernels and CUDA)."""
    got = np.asarray(got, dtype=float).reshape(-1, 8) if np.size(got) else np.empty((0, 8))
    want = np.asarray(want, dtype=float).reshape(-1, 8) if np.size(want) else np.empty((0, 8))
    assert got.shape == want.shape, "frame %d: %s vs %s rows" % (frame, got.shape, want.shape)
    if got.size == 0:
        return
    assert np.array_equal(got[:, 4], want[:, 4]), "frame %d: track ids differ" % frame
    assert np.array_equal(got[:, 7], want[:, 7]), "frame %d: det_ind differ" % frame
    if conf_rtol:
        assert np.array_equal(got[:, 6], want[:, 6]), "frame %d: cls differ" % frame
        np.testing.assert_allclose(got[:, 5], want[:, 5], rtol=conf_rtol, atol=0, err_msg="frame %d conf" % frame)
    else:
        assert np.array_equal(got[:, 5:7], want[:, 5:7]), "frame %d: conf/cls differ" % frame
    np.testing.assert_allclose(got[:, :4], want[:, :4], rtol=box_rtol, atol=1e-6,
                               err_msg="frame %d boxes" % frame)
