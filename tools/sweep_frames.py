"""Developer sweep: OC-SORT throughput of bm_ocsort_step_frames over (groups, frames per call)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as B
from boxmot_b200.engine import OcSortEngine
S, K, W, F = 1024, 100, 10, 170
dets, nd = B.gen_inputs(0, S, K, F, 128, 16)
eng = OcSortEngine(S, track_cap=256, det_cap=128, **B.OCSORT_YAML)
d_dets, d_nd = torch.as_tensor(dets).cuda(), torch.as_tensor(nd).cuda()
flush = torch.empty(2 * B.L2_BYTES, dtype=torch.uint8, device="cuda")
for G in (1, 2, 4, 8, 16):
    for C in (1, 4, 8, 16, 32):
        eng.set_stream_groups(G)
        eng.reset()
        for f in range(W):
            eng.step(d_dets[f], d_nd[f])
        c0 = eng.counters()
        out = torch.zeros((C, S, eng.out_rows, 8), dtype=torch.float64, device="cuda")
        nout = torch.zeros((C, S), dtype=torch.int32, device="cuda")
        tot = 0.0
        n = 0
        for f0 in range(W, F - C + 1, C):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            eng.step_frames(d_dets[f0:f0 + C], d_nd[f0:f0 + C], out, nout)
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
            n += C
        c1 = eng.counters()
        eng.status()
        print("groups %2d frames/call %2d: %.4f ms/frame  %.1f M track-updates/s" % (
            G, C, tot / n, (c1["track_updates"] - c0["track_updates"]) / tot / 1e3), flush=True)
