"""Debug: nn cosine fast path vs fixture, element by element."""
import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from boxmot_b200 import _lib
from scipy.spatial.distance import cdist
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "ref_strongsort_units.npz"))
bank, nb, feats = z["bank"], z["nb"].astype(np.int32), z["feats"]
T, B, E = bank.shape
D = feats.shape[0]
print("dtypes", bank.dtype, feats.dtype, bank.shape, feats.shape)
d_bank, d_nb, d_f = torch.as_tensor(bank).cuda(), torch.as_tensor(nb).cuda(), torch.as_tensor(feats).cuda()
for rep in range(2):
    d_out = torch.full((T, D), -7.0, dtype=torch.float32, device="cuda")
    d_work = torch.zeros((max(B, D) + T * B + D + 16,), dtype=torch.int32, device="cuda")
    rc = _lib.lib().bm_nn_cosine_distance(d_bank.data_ptr(), d_nb.data_ptr(), T, B, d_f.data_ptr(), D, E,
                                          d_out.data_ptr(), d_work.data_ptr(), None)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy().astype(np.float64)
    err = np.abs(got - z["nn"])
    print("rep", rep, "rc", rc, "max err", err.max(), "n bad", (err > 2e-6).sum())
    for t in range(T):
        print(" t", t, "nb", nb[t], "max err %.3g" % err[t].max(), "bad cols", np.nonzero(err[t] > 2e-6)[0][:12],
              "got", got[t, :3], "want", z["nn"][t, :3])
# the same data through bm_cosine_distance (raw = 0), per track
for t in (0, 1, 7, 11):
    if nb[t] == 0:
        continue
    a = feats.astype(np.float32); b = bank[t, : nb[t]].astype(np.float32)
    d_a, d_b = torch.as_tensor(a).cuda().contiguous(), torch.as_tensor(b).cuda().contiguous()
    out = torch.full((a.shape[0], b.shape[0]), -7.0, dtype=torch.float32, device="cuda")
    work = torch.zeros((2 * (a.shape[0] + b.shape[0]),), dtype=torch.int32, device="cuda")
    p = lambda x: ctypes.c_void_p(x.data_ptr())
    rc = _lib.lib().bm_cosine_distance(p(d_a), a.shape[0], p(d_b), b.shape[0], E, p(out), p(work), 0, None)
    torch.cuda.synchronize()
    want = np.maximum(0, cdist(a.astype(np.float64), b.astype(np.float64), "cosine"))
    print("cosine t", t, "rc", rc, "max err", np.abs(out.cpu().numpy() - want).max())
