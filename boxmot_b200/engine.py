"""Stream-batched device engines (host side).  PyTorch supplies device memory and the
CUDA stream; all arithmetic happens in libboxmot_b200.so (see include/boxmot_b200.h)."""
import ctypes

import numpy as np
import torch

from boxmot_b200 import _lib

ERR_BITS = {1: "track capacity exceeded", 2: "detection capacity exceeded",
            4: "cost matrix larger than scratch", 8: "non-finite costs reached the assignment solver"}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class OcSortEngine:
    """S independent OC-SORT trackers advanced by one kernel launch per frame.

    Parameters are those of the reference ``OcSort.__init__``
    (boxmot/trackers/ocsort/ocsort.py:213-227); ``track_cap`` / ``det_cap`` bound the
    live tracks / detections per stream (overflow raises, it never truncates)."""

    def __init__(self, n_streams, track_cap=256, det_cap=128, device="cuda:0", min_conf=0.1,
                 det_thresh=0.2, max_age=30, min_hits=3, asso_threshold=0.3, delta_t=3,
                 asso_func="iou", inertia=0.2, use_byte=False, Q_xy_scaling=0.01,
                 Q_s_scaling=0.0001, img_wh=(1920, 1080), smem_cost_bytes=None):
        if not torch.cuda.is_available():
            raise RuntimeError("OcSortEngine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.S, self.Tcap, self.Dcap = int(n_streams), int(track_cap), int(det_cap)
        if smem_cost_bytes is None:
            smem_cost_bytes = 0     # second-round matrices live in the L2-resident scratch
        cfg = _lib.OcsortConfig(det_thresh, min_conf, asso_threshold, inertia, Q_xy_scaling,
                                Q_s_scaling, int(max_age), int(min_hits), int(delta_t),
                                _lib.ASSO_FUNCS[asso_func], int(bool(use_byte)),
                                int(img_wh[0]), int(img_wh[1]))
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_create(ctypes.byref(cfg), self.S, self.Tcap, self.Dcap,
                                                   int(smem_cost_bytes), ctypes.byref(self._h)),
                       "bm_ocsort_create")
            torch.cuda.synchronize(self.device)
        self.out = torch.zeros((self.S, self.Tcap, 8), dtype=torch.float64, device=self.device)
        self.nout = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
        self.launches = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            _lib.lib().bm_ocsort_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_reset(self._h, _stream_ptr(self.device)), "bm_ocsort_reset")

    def step(self, dets, ndets):
        """dets: float32 cuda tensor [S, det_cap, 6]; ndets: int32 cuda tensor [S].
        Asynchronous; returns the engine-owned (out [S,track_cap,8] f64, nout [S] i32)."""
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_step(self._h, _ptr(dets), _ptr(ndets), _ptr(self.out),
                                                 _ptr(self.nout), _stream_ptr(self.device)),
                       "bm_ocsort_step")
        self.launches += 1
        return self.out, self.nout

    def status(self, raise_on_error=True):
        err = ctypes.c_int(0)
        upd = ctypes.c_ulonglong(0)
        laps = ctypes.c_ulonglong(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_status(self._h, ctypes.byref(err), ctypes.byref(upd),
                                                   ctypes.byref(laps), _stream_ptr(self.device)),
                       "bm_ocsort_status")
        if raise_on_error and err.value:
            msgs = [m for b, m in ERR_BITS.items() if err.value & b]
            raise RuntimeError("OC-SORT engine error: " + "; ".join(msgs))
        return {"err": err.value, "track_updates": upd.value, "lap_solves": laps.value}

    def counters(self):
        """Cumulative {track_updates, lap_solves, matches, births} over all streams."""
        buf = (ctypes.c_ulonglong * 4)()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_counters(self._h, ctypes.byref(buf),
                                                     _stream_ptr(self.device)), "bm_ocsort_counters")
        return dict(zip(("track_updates", "lap_solves", "matches", "births"), [int(v) for v in buf]))

    def peek(self, stream_index):
        n = ctypes.c_int(0)
        x = np.zeros((self.Tcap, 7))
        P = np.zeros((self.Tcap, 49))
        ids = np.zeros((self.Tcap,), dtype=np.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_peek(self._h, int(stream_index), ctypes.byref(n),
                                                 x.ctypes.data, P.ctypes.data, ids.ctypes.data,
                                                 _stream_ptr(self.device)), "bm_ocsort_peek")
        k = n.value
        return x[:k], P[:k].reshape(k, 7, 7), ids[:k]

    # -- host-buffer convenience (the reference-facing path: H2D, step, D2H) -------------
    def step_host(self, dets_per_stream):
        """dets_per_stream: list of S arrays (N_s, 6).  Returns list of (M_s, 8) arrays."""
        if not hasattr(self, "_h_dets"):
            self._h_dets = torch.zeros((self.S, self.Dcap, 6), dtype=torch.float32).pin_memory()
            self._h_nd = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
            self._d_dets = torch.zeros((self.S, self.Dcap, 6), dtype=torch.float32, device=self.device)
            self._d_nd = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
            self._h_out = torch.zeros((self.S, self.Tcap, 8), dtype=torch.float64).pin_memory()
            self._h_nout = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
        hd = self._h_dets.numpy()
        hn = self._h_nd.numpy()
        for s, d in enumerate(dets_per_stream):
            n = 0 if d is None else len(d)
            if n > self.Dcap:
                raise RuntimeError("stream %d: %d detections exceed det_cap=%d" % (s, n, self.Dcap))
            if n:
                hd[s, :n] = d
            hn[s] = n
        self._d_dets.copy_(self._h_dets, non_blocking=True)
        self._d_nd.copy_(self._h_nd, non_blocking=True)
        out, nout = self.step(self._d_dets, self._d_nd)
        self._h_out.copy_(out, non_blocking=True)
        self._h_nout.copy_(nout, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        ho, hn2 = self._h_out.numpy(), self._h_nout.numpy()
        return [ho[s, : hn2[s]].copy() for s in range(self.S)]
