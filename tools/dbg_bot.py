import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import BOTSORT_YAML
from boxmot_b200.engine import BotSortEngine
from boxmot_b200.synth import make_batch
from oracle.botsort import BotSortOracle, embedding_distance
from oracle import assoc
S, K, F, D, E = 32, 30, 50, 64, 64
dets, nd, embs = make_batch(S, K, F, D, emb_dim=E, first_stream=500)
eng = BotSortEngine(S, track_cap=128, det_cap=D, emb_dim=E, **BOTSORT_YAML)
oracles = [BotSortOracle(**BOTSORT_YAML) for _ in range(S)]
d_dets = torch.as_tensor(dets.astype(np.float32)).cuda(); d_nd = torch.as_tensor(nd).cuda(); d_embs = torch.as_tensor(embs).cuda()
for f in range(8):
    out, nout = eng.step(d_dets[f], d_nd[f], d_embs[f])
    out, nout = out.cpu().numpy(), nout.cpu().numpy()
    for s in range(S):
        n = nd[f, s]
        o = oracles[s]
        want = o.update(dets[f, s, :n], embs[f, s, :n].copy())
        got = out[s, :nout[s]]
        if got.shape != want.shape or not np.array_equal(got[:, [4, 7]], want[:, [4, 7]]):
            print("MISMATCH frame", f, "stream", s)
            print(" got ", got[:, [4, 7]].T)
            print(" want", want[:, [4, 7]].T)
            bad = np.nonzero(got[:, 7] != want[:, 7])[0] if got.shape == want.shape else []
            for b in bad:
                print("  id", want[b, 4], "conf got", got[b, 5], "want", want[b, 5], "boxes", got[b, :4], want[b, :4])
            sys.exit(0)
print("no mismatch in 8 frames")
