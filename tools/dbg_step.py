import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from boxmot_b200.engine import OcSortEngine
from boxmot_b200.synth import make_batch
kw = dict(min_conf=0.1, det_thresh=0.6, max_age=30, min_hits=3, asso_threshold=0.3, delta_t=3, asso_func="iou", inertia=0.1, use_byte=False, Q_xy_scaling=0.01, Q_s_scaling=0.0001)
S,K,F,D=[int(x) for x in (sys.argv[1:5] or [8,30,25,64])]
TC=int(sys.argv[5]) if len(sys.argv)>5 else 128
dets, nd = make_batch(S,K,F,D, first_stream=int(os.environ.get("FS","7000")))
eng = OcSortEngine(S, track_cap=TC, det_cap=D, **kw)
dd=torch.as_tensor(dets.astype(np.float32)).cuda(); dn=torch.as_tensor(nd).cuda()
for f in range(F):
    eng.step(dd[f], dn[f]); torch.cuda.synchronize(); print("frame", f, "ok", eng.status(False))
