import sys, os
import numpy as np, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from boxmot_b200.engine import StrongSortEngine
from boxmot_b200.synth import make_batch
from conftest import STRONGSORT_YAML
from oracle.strongsort_tracker import StrongSortOracle
from oracle import strongsort as so
NS, KO, FR, E, DC = 2, 100, 30, 768, 128
kw = dict(STRONGSORT_YAML, nn_budget=50)
dets, nd, embs = make_batch(96, KO, FR, DC, emb_dim=E)
sel = [0, 95]
dets, nd, embs = dets[:, sel].astype(np.float32), nd[:, sel], embs[:, sel]
eng = StrongSortEngine(NS, track_cap=1024, det_cap=DC, pool_cap=4096, emb_dim=E, **kw)
orcs = [StrongSortOracle(**kw) for _ in sel]
LOG = {}
_orig = StrongSortOracle._enhanced_match
def _em(self, dets):
    m, ut, ud = _orig(self, dets)
    LOG[id(self)] = (list(m), [self.tracks[t].id for t, _ in m], [dets[d]["det_ind"] for _, d in m],
                     [(self.tracks[t].id, int(self.tracks[t].a[so.TRK_TSU])) for t in range(len(self.tracks))])
    return m, ut, ud
StrongSortOracle._enhanced_match = _em
LAST = {}
_omcm = StrongSortOracle._min_cost_matching
def _mcm(self, cost, tab, dets, tis, dis, max_distance):
    q = np.array([dets[i]["quality"] for i in dis]); c = np.array([dets[i]["conf"] for i in dis]); k = np.array([dets[i]["cls"] for i in dis])
    cc = so.enhance_cost_matrix(cost, tab, q, c, k)
    LAST[id(self)] = (cc.copy(), list(tis), list(dis), max_distance, np.array(cost))
    return _omcm(self, cost, tab, dets, tis, dis, max_distance)
StrongSortOracle._min_cost_matching = _mcm
d_dets, d_nd = torch.as_tensor(np.ascontiguousarray(dets)).cuda(), torch.as_tensor(np.ascontiguousarray(nd)).cuda()
for f in range(FR):
    out, nout = eng.step(d_dets[f], d_nd[f], torch.as_tensor(np.ascontiguousarray(embs[f])).cuda())
    out, nout = out.cpu().numpy(), nout.cpu().numpy()
    for s in range(NS):
        n = nd[f, s]
        want = orcs[s].update(dets[f, s, :n].astype(np.float64), embs[f, s, :n].astype(np.float64))
        got = out[s, :nout[s]]
        x, P, tab, ids, nb, nl = eng.peek(s)
        o = orcs[s]
        oid = np.array([t.id for t in o.tracks])
        if got.shape != want.shape or not np.array_equal(got[:, 4:], want[:, 4:]):
            print("frame", f, "stream", s, "rows differ", got.shape, want.shape)
            k = min(len(got), len(want))
            d = np.nonzero(np.any(got[:k, 4:] != want[:k, 4:], axis=1))[0]
            print(d[:5], got[d[0]] if len(d) else None, want[d[0]] if len(d) else None)
            m, tids, dinds, tsu = LOG[id(orcs[s])]
            tid = int(want[d[0], 4])
            print("oracle matches of track", tid, [(i, m[i], dinds[i]) for i in range(len(m)) if tids[i] == tid])
            print("tsu before match", [x for x in tsu if x[0] == tid], "n matches", len(m), "n tracks", len(tsu))
            import ctypes
            from boxmot_b200 import _lib
            from scipy.optimize import linear_sum_assignment
            Tc, Dc = eng.Tcap, eng.Dcap
            wk = np.zeros(3 * Tc + 2 * Tc + 2 * Tc + Dc + 2 * Tc + Dc + Dc, np.int32)
            cost = np.zeros((Tc, Dc))
            _lib.check(_lib.lib().bm_strongsort_debug_copy(eng._h, 0, s, wk.ctypes.data, wk.nbytes, None), "dbg")
            _lib.check(_lib.lib().bm_strongsort_debug_copy(eng._h, 1, s, cost.ctypes.data, cost.nbytes, None), "dbg")
            o0 = 3 * Tc
            rows3 = wk[o0:o0 + 2 * Tc]; mcol = wk[o0 + 2 * Tc:o0 + 4 * Tc]; cols = wk[o0 + 4 * Tc:o0 + 4 * Tc + Dc]
            ntr = len(tsu)
            tsu_by_pos = np.array([x[1] for x in tsu])
            cand = [p for p in range(ntr)] + [p for p in range(ntr) if tsu_by_pos[p] == 1]
            print("gpu rows3 head/tail", rows3[:5], rows3[len(cand) - 5:len(cand) + 2], "expected n cand", len(cand))
            print("gpu mcol for pos 355:", [(r, mcol[r]) for r in range(len(cand)) if rows3[r] == 355])
            dt = np.zeros((Dc, 13))
            _lib.check(_lib.lib().bm_strongsort_debug_copy(eng._h, 2, s, dt.ctypes.data, dt.nbytes, None), "dbg")
            nn = int(nd[f, s])
            conf = dt[:, 8]
            remd = [k for k in range(nn) if conf[k] >= 0.3]          # min_conf 0.1 keeps all; stage 1/2 matched nothing
            cst = 0.7 + 1e-5
            M = np.full((len(cand), len(remd)), cst)
            for r, p in enumerate(cand):
                if tsu_by_pos[p] <= 1:
                    row = cost[p, remd]
                    M[r] = np.where(row > 0.7, cst, row)
            rows_, cols_ = linear_sum_assignment(M)
            sc = -np.ones(len(cand), int); sc[rows_] = cols_
            gpu = np.array([remd.index(mcol[r]) if mcol[r] >= 0 else -1 for r in range(len(cand))])
            acc = np.array([sc[r] if sc[r] >= 0 and M[r, sc[r]] <= 0.7 else -1 for r in range(len(cand))])
            print("scipy on the GPU matrix == GPU accepted pairs:", np.array_equal(acc, gpu), "diff rows", np.nonzero(acc != gpu)[0][:10])
            print("scipy rows 355/449 ->", [remd[c] if c >= 0 else -1 for c in (acc[355], acc[449])])
            cc, tis, dis, md, raw = LAST[id(orcs[s])]
            print("oracle last mcm: shape", cc.shape, "max_distance", md, "tis==cand", tis == cand, "dis==remd", dis == remd)
            OM = np.where(cc > md, md + 1e-5, cc)
            dd = np.abs(OM - M)
            print("max |oracle - gpu| matrix diff", dd.max(), "n differing entries", int((OM != M).sum()))
            bad = np.argwhere(OM != M)[:8]
            for r, c in bad:
                print("  row", r, "pos", cand[r], "det", remd[c], "oracle", repr(OM[r, c]), "gpu", repr(M[r, c]), "raw iou cost", repr(raw[r, c]))
            sys.exit()
        if len(oid) != len(ids) or not np.array_equal(oid, ids):
            print("frame", f, "stream", s, "track list differs", len(oid), len(ids)); sys.exit()
        otab = np.array([t.a for t in o.tracks]); otab[:, so.TS_MISSED] = tab[:, so.TS_MISSED]
        dt = np.abs(otab - tab).max(axis=0)
        onb = np.array([len(o.bank.get(t.id, [])) for t in o.tracks])
        if dt.max() > 1e-9 or not np.array_equal(onb, nb):
            print("frame", f, "stream", s, "state differs cols", np.nonzero(dt > 1e-9)[0], dt.max(), "nbank", np.array_equal(onb, nb)); sys.exit()
print("all equal")
