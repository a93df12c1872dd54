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
from boxmot_b200.synth import make_stream  # noqa: E402

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
    main()
