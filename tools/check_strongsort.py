"""Developer check: StrongSORT engine vs the committed reference fixtures and the numpy oracle."""
import os
import sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from boxmot_b200.engine import StrongSortEngine
from boxmot_b200.synth import strongsort_stream
from oracle.strongsort_tracker import StrongSortOracle
from oracle import strongsort as so

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
CASES = [
    ("synth_k20_e32", 60, 20, 60, 32, False, {}),
    ("synth_k30_e64_b20", 61, 30, 70, 64, False, {"nn_budget": 20}),
    ("synth_k40_e32_cmc", 62, 40, 50, 32, True, {"nn_budget": 50}),
    ("synth_k25_e96_age8", 63, 25, 80, 96, False, {"max_age": 8, "nn_budget": 12}),
    ("synth_k60_e32_crowd", 64, 60, 40, 32, False, {"nn_budget": 30, "max_age": 10}),
    ("synth_k15_e32_ninit3", 65, 15, 90, 32, True, {"n_init": 3, "max_age": 6, "nn_budget": 8, "min_conf": 0.3}),
    ("synth_k50_e128_hi", 66, 50, 40, 128, False, {"conf_thresh_high": 0.5, "conf_thresh_low": 0.2, "max_cos_dist": 0.3}),
]


def inputs(sid, k, f, e, crowd):
    return strongsort_stream(sid, k, f, e, crowd)


for name, sid, k, f, e, use_warps, kw in CASES:
    pick = [a for a in sys.argv[1:] if not a.startswith("--")]
    if pick and not any(a in name for a in pick):
        continue
    z = np.load(os.path.join(GOLD, "ref_strongsort_%s.npz" % name))
    rows = z["rows"]
    warps = z["warps"] if use_warps else None
    fr, em = inputs(sid, k, f, e, "crowd" in name)
    eng = StrongSortEngine(1, track_cap=1024, det_cap=128, pool_cap=4096, emb_dim=e, **kw)
    orc = StrongSortOracle(**kw) if "--oracle" in sys.argv else None
    bad = None
    for i, d in enumerate(fr):
        w = None if warps is None else [warps[i]]
        got = eng.step_host([d], [em[i]], w)[0]
        want = rows[rows[:, 0] == i][:, 1:]
        if orc is not None:
            ow = orc.update(d, em[i], None if warps is None else warps[i])
            assert ow.shape == want.shape
            x, P, tab, ids, nb, nl = eng.peek(0)
            oid = np.array([t.id for t in orc.tracks])
            if len(oid) != len(ids) or not np.array_equal(oid, ids):
                print("frame", i, "track list differs", len(oid), len(ids)); bad = i; break
            otab = np.array([t.a for t in orc.tracks])
            omean = np.array([t.mean for t in orc.tracks])
            onb = np.array([len(orc.bank.get(t.id, [])) for t in orc.tracks])
            if len(oid):
                dt = np.abs(otab - tab).max(axis=0)
                dt[so.TS_MISSED] = 0        # missed_detections is never read (a duplicated IoU row counts twice)
                if dt.max() > 1e-9 or np.abs(omean - x).max() > 1e-7 or not np.array_equal(onb, nb):
                    print("frame", i, "state differs: tab cols", np.nonzero(dt > 1e-9)[0], "mean", np.abs(omean - x).max(),
                          "nbank eq", np.array_equal(onb, nb)); bad = i; break
        ok = got.shape == want.shape and np.array_equal(got[:, 4:], want[:, 4:]) and \
            np.allclose(got[:, :4], want[:, :4], rtol=1e-6, atol=1e-6)
        if not ok:
            bad = i
            print("frame", i, got.shape, want.shape)
            n = min(len(got), len(want))
            dd = np.nonzero(np.any(got[:n, 4:] != want[:n, 4:], axis=1))[0]
            if len(dd):
                print(got[dd[0]], want[dd[0]])
            break
    st = eng.counters(raise_on_error=False)
    print(name, "OK" if bad is None else "MISMATCH at frame %d" % bad, st)
    eng.close()
