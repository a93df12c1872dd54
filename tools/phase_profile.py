"""Per-phase clock64 breakdown of ocs_assoc_kernel (debug): mean / p99 / max cycles over streams."""
import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from boxmot_b200 import _lib
from boxmot_b200.engine import OcSortEngine
from boxmot_b200.synth import make_batch
kw = dict(min_conf=0.1, det_thresh=0.6, max_age=30, min_hits=3, asso_threshold=0.3, delta_t=3, asso_func="iou",
          inertia=0.1, use_byte=False, Q_xy_scaling=0.01, Q_s_scaling=0.0001)
S, F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 60
dets, nd = make_batch(S, 100, F, 128)
eng = OcSortEngine(S, track_cap=256, det_cap=128, **kw)
dd = torch.as_tensor(dets.astype(np.float32)).cuda(); dn = torch.as_tensor(nd).cuda()
st = torch.zeros((S, 8), dtype=torch.int64, device="cuda")
_lib.check(_lib.lib().bm_ocsort_debug_profile(eng._h, ctypes.c_void_p(st.data_ptr())), "prof")
names = ["staging+screen", "row scan", "first-round solve", "lists", "byte", "ocr", "births"]
for f in range(F):
    eng.step(dd[f], dn[f])
    if f in (1, 2, F // 2, F - 1):
        torch.cuda.synchronize()
        a = st.cpu().numpy().astype(np.float64)
        d = np.diff(a[:, :7], axis=1)
        tot = a[:, 6] - a[:, 0]
        print("frame %d: CTA lifetime mean %.0f p99 %.0f max %.0f cycles; kernel span %.0f cycles" % (
            f, tot.mean(), np.percentile(tot, 99), tot.max(), a[:, 6].max() - a[:, 0].min()))
        for i, n in enumerate(names[:6]):
            print("   %-18s mean %8.0f  p99 %8.0f  max %8.0f" % (n, d[:, i].mean(), np.percentile(d[:, i], 99), d[:, i].max()))
        worst = np.argsort(-tot)[:8]
        print("   slowest CTAs (stream: phases):")
        for w in worst:
            print("     %4d: %s  nd=%d" % (w, " ".join("%7.0f" % x for x in d[w, :6]), nd[f, w]))
        print("   lifetime histogram (k cycles):", np.histogram(tot / 1e3, bins=[0, 80, 120, 160, 200, 240, 280, 320, 400, 1000])[0])
eng.status()
