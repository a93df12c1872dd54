"""Developer check: capacity the StrongSORT engine needs on the configs[3] streams (K objects/frame)."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from boxmot_b200.engine import StrongSortEngine
from boxmot_b200.synth import make_batch
S, K, F, E = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
B = int(sys.argv[5]) if len(sys.argv) > 5 else 50
dets, nd, embs = make_batch(S, K, F, 128 if K <= 100 else 2 * K, emb_dim=E)
D = dets.shape[2]
eng = StrongSortEngine(S, track_cap=4096, det_cap=D, pool_cap=16384, emb_dim=E, nn_budget=B)
d_dets, d_nd = torch.as_tensor(dets.astype(np.float32)).cuda(), torch.as_tensor(nd).cuda()
eng.time_kernels(True)
for f in range(F):
    d_e = torch.as_tensor(embs[f]).cuda()
    eng.step(d_dets[f], d_nd[f], d_e)
    if f % 10 == 9 or f == F - 1:
        print(f, eng.usage(), ["%.3f" % x for x in eng.kernel_times()], eng.counters(False))

import ctypes
from boxmot_b200 import _lib
acc = np.zeros(7)
for s_ in range(0, S, max(1, S // 16)):
    st = np.zeros(10, np.int64)
    _lib.check(_lib.lib().bm_strongsort_debug_copy(eng._h, 3, s_, st.ctypes.data, st.nbytes, None), "dbg")
    acc += np.diff(st[:8])
    print("  stream", s_, "partial_fit pass 1 cycles", st[8] - st[4], "targets with work (ss_bank_kernel)", st[9])
print("finalize phases (cycles, mean of sampled streams): mark_missed, id recovery, births, deletions, partial_fit pass 1, outputs:",
      (acc / len(range(0, S, max(1, S // 16)))).astype(int)[:6])
