"""Advance S synthetic streams by F frames through an engine (driver for ncu captures).
usage: run_frames.py [S=1024] [F=45] [K=100] [tracker=ocsort|bytetrack|botsort|strongsort|deepocsort|boosttrack] [E=512]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from boxmot_b200.engine import ByteTrackEngine, OcSortEngine
from boxmot_b200.synth import make_batch
S = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
F = int(sys.argv[2]) if len(sys.argv) > 2 else 45
K = int(sys.argv[3]) if len(sys.argv) > 3 else 100
tracker = sys.argv[4] if len(sys.argv) > 4 else "ocsort"
embs = None
if tracker == "botsort":
    from boxmot_b200.engine import BotSortEngine
    E = int(sys.argv[5]) if len(sys.argv) > 5 else 512
    dets, nd, embs = make_batch(S, K, F, 256, emb_dim=E)
    eng = BotSortEngine(S, track_cap=512, det_cap=256, emb_dim=E, track_high_thresh=0.6, track_low_thresh=0.1,
                        new_track_thresh=0.7, track_buffer=30, match_thresh=0.8, proximity_thresh=0.5,
                        appearance_thresh=0.25, frame_rate=30, fuse_first_associate=False, with_reid=True)
elif tracker == "strongsort":
    import bench
    from boxmot_b200.engine import StrongSortEngine
    E = int(sys.argv[5]) if len(sys.argv) > 5 else 768
    dets, nd, embs = make_batch(S, K, F, 128, emb_dim=E)
    eng = StrongSortEngine(S, track_cap=2048, det_cap=128, pool_cap=8192, emb_dim=E, **bench.STRONGSORT_ARGS)
elif tracker in ("deepocsort", "boosttrack"):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from boxmot_b200.engine import BoostTrackEngine, DeepOcSortEngine
    E = int(sys.argv[5]) if len(sys.argv) > 5 else 512
    dets, nd, embs = make_batch(S, K, F, 128, emb_dim=E)
    if tracker == "deepocsort":
        eng = DeepOcSortEngine(S, track_cap=256, det_cap=128, emb_dim=E, **bench.DEEPOCSORT_YAML)
    else:
        eng = BoostTrackEngine(S, track_cap=256, det_cap=128, emb_dim=E, **bench.BOOSTTRACK_YAML)
else:
    dets, nd = make_batch(S, K, F, 128)
if tracker in ("botsort", "deepocsort", "boosttrack", "strongsort"):
    pass
elif tracker == "bytetrack":
    eng = ByteTrackEngine(S, track_cap=256, det_cap=128, min_conf=0.1, track_thresh=0.6, match_thresh=0.9,
                          track_buffer=30, frame_rate=30)
else:
    eng = OcSortEngine(S, track_cap=256, det_cap=128, min_conf=0.1, det_thresh=0.6, max_age=30, min_hits=3,
                       asso_threshold=0.3, delta_t=3, asso_func="iou", inertia=0.1, use_byte=False,
                       Q_xy_scaling=0.01, Q_s_scaling=0.0001)
dd = torch.as_tensor(dets.astype(np.float32)).cuda(); dn = torch.as_tensor(nd).cuda()
de = None if embs is None else torch.as_tensor(embs).cuda()
for f in range(F):
    if de is None:
        eng.step(dd[f], dn[f])
    else:
        eng.step(dd[f], dn[f], de[f])
torch.cuda.synchronize()
print(eng.status(), eng.counters())
