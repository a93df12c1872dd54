"""Stream-batched device engines (host side).  PyTorch supplies device memory and the
CUDA stream; all arithmetic happens in libboxmot_b200.so (see include/boxmot_b200.h)."""
import ctypes

import numpy as np
import torch

from boxmot_b200 import _lib

ERR_BITS = {1: "track capacity exceeded", 2: "detection capacity exceeded",
            4: "cost matrix larger than scratch", 8: "non-finite costs reached the assignment solver",
            16: "a track saw more distinct classes than the class-vote history holds",
            32: "camera-motion warp the engine could not apply (reserved: rotating warps are handled by the dense-covariance path)"}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _HostStepMixin:
    # -- host-buffer convenience (the reference-facing path: H2D, step, D2H) -------------
    def step_host(self, dets_per_stream):
        """dets_per_stream: list of S arrays (N_s, 6).  Returns list of (M_s, 8) arrays."""
        if not hasattr(self, "_h_dets"):
            self._h_dets = torch.zeros((self.S, self.Dcap, 6), dtype=torch.float32).pin_memory()
            self._h_nd = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
            self._d_dets = torch.zeros((self.S, self.Dcap, 6), dtype=torch.float32, device=self.device)
            self._d_nd = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
            self._h_out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64).pin_memory()
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


def _step_classes_host(eng, class_dets, class_embs=None, class_warps=None):
    """per_class stepping of a ONE-stream engine whose lost / removed lists (ByteTrack, BoT-SORT) or whole track
    pool (BoostTrack) are shared by the classes: ``class_dets[c]`` (N_c, 6) are the detections of class id c.
    All classes are uploaded at once; per class ``eng.select_class`` then ``eng.step`` on that class's slice are
    enqueued without a host synchronisation; one download.  Returns the list of (M_c, 8) row arrays."""
    assert eng.S == 1, "per_class stepping drives a single-stream engine"
    C = len(class_dets)
    E = getattr(eng, "E", 0) if getattr(eng, "with_reid", False) else 0
    key = "_pc_bufs"
    if getattr(eng, key, None) is None or eng._pc_bufs[0] != C:
        h = dict(dets=torch.zeros((C, eng.Dcap, 6), dtype=torch.float32).pin_memory(),
                 nd=torch.zeros((C,), dtype=torch.int32).pin_memory(),
                 embs=torch.zeros((C, eng.Dcap, max(E, 1)), dtype=torch.float32).pin_memory(),
                 warps=torch.zeros((C, 2, 3), dtype=torch.float64).pin_memory(),
                 out=torch.zeros((C, eng.out_rows, 8), dtype=torch.float64).pin_memory(),
                 nout=torch.zeros((C,), dtype=torch.int32).pin_memory())
        d = {k: torch.zeros_like(v, device=eng.device) for k, v in h.items()}
        eng._pc_bufs = (C, h, d)
    _, h, d = eng._pc_bufs
    hd, hn, he, hw = h["dets"].numpy(), h["nd"].numpy(), h["embs"].numpy(), h["warps"].numpy()
    for c, x in enumerate(class_dets):
        n = 0 if x is None else len(x)
        if n > eng.Dcap:
            raise RuntimeError("class %d: %d detections exceed det_cap=%d" % (c, n, eng.Dcap))
        if n:
            hd[c, :n] = x
            if E:
                he[c, :n] = class_embs[c]
        hn[c] = n
        hw[c] = np.eye(2, 3) if class_warps is None else class_warps[c]
    for k in ("dets", "nd") + (("embs",) if E else ()) + (("warps",) if class_warps is not None else ()):
        d[k].copy_(h[k], non_blocking=True)
    for c in range(C):
        eng.select_class(C, c, c == 0)
        args = [d["dets"][c:c + 1], d["nd"][c:c + 1]]
        if hasattr(eng, "with_reid"):
            args += [d["embs"][c:c + 1] if E else None, d["warps"][c:c + 1] if class_warps is not None else None]
        eng.step(*args, out=d["out"][c:c + 1], nout=d["nout"][c:c + 1])
    h["out"].copy_(d["out"], non_blocking=True)
    h["nout"].copy_(d["nout"], non_blocking=True)
    torch.cuda.current_stream(eng.device).synchronize()
    ho, hn2 = h["out"].numpy(), h["nout"].numpy()
    return [ho[c, : hn2[c]].copy() for c in range(C)]


class OcSortEngine(_HostStepMixin):
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
        self.out_rows = min(self.Tcap, self.Dcap)
        self.out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
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

    def step(self, dets, ndets, out=None, nout=None):
        """dets: float32 cuda tensor [S, det_cap, 6]; ndets: int32 cuda tensor [S].
        Asynchronous; returns (out [S,out_rows,8] f64, nout [S] i32) -- the engine-owned
        buffers unless the caller passes its own (e.g. to double-buffer the D2H copy)."""
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        out = self.out if out is None else out
        nout = self.nout if nout is None else nout
        assert out.dtype == torch.float64 and tuple(out.shape) == (self.S, self.out_rows, 8) and out.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_step(self._h, _ptr(dets), _ptr(ndets), _ptr(out),
                                                 _ptr(nout), _stream_ptr(self.device)),
                       "bm_ocsort_step")
        self.launches += 4      # predict, assoc, update, finalize
        return out, nout

    def step_frames(self, dets, ndets, out=None, nout=None):
        """F consecutive frames in one call: dets [F,S,det_cap,6] float32, ndets [F,S] int32 -> out
        [F,S,out_rows,8], nout [F,S].  Same results as F ``step`` calls; with stream groups the groups run
        all their frames back to back and are joined once (bm_ocsort_step_frames)."""
        F = int(dets.shape[0])
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (F, self.S, self.Dcap, 6) and tuple(ndets.shape) == (F, self.S)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.is_contiguous()
        if out is None:
            out = torch.zeros((F, self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
            nout = torch.zeros((F, self.S), dtype=torch.int32, device=self.device)
        assert tuple(out.shape) == (F, self.S, self.out_rows, 8) and out.is_contiguous() and nout.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_step_frames(self._h, _ptr(dets), _ptr(ndets), _ptr(out), _ptr(nout), F,
                                                        _stream_ptr(self.device)), "bm_ocsort_step_frames")
        self.launches += 4 * F
        return out, nout

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

    def set_stream_groups(self, groups):
        """Step the streams as ``groups`` independent groups on CUDA streams of their own (forked
        from / joined into the current stream inside ``step``); returns the number in use."""
        with torch.cuda.device(self.device):
            n = _lib.lib().bm_ocsort_set_groups(self._h, int(groups))
        if n < 0:
            _lib.check(n, "bm_ocsort_set_groups")
        self.groups = n
        return n

    def stream_info(self):
        """(live tracks [S], tracks created so far [S]) as int32 arrays (synchronises)."""
        n = np.zeros((self.S,), dtype=np.int32)
        nid = np.zeros((self.S,), dtype=np.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_ocsort_stream_info(self._h, n.ctypes.data, nid.ctypes.data,
                                                        _stream_ptr(self.device)), "bm_ocsort_stream_info")
        return n, nid

    def time_kernels(self, enable=True):
        """Bracket the four launches of every following step with CUDA events."""
        _lib.check(_lib.lib().bm_ocsort_time_kernels(self._h, int(bool(enable))), "bm_ocsort_time_kernels")

    def kernel_times(self):
        """(predict, assoc, update, finalize) milliseconds of the last timed step (synchronises)."""
        buf = (ctypes.c_float * 4)()
        _lib.check(_lib.lib().bm_ocsort_kernel_times(self._h, ctypes.byref(buf)), "bm_ocsort_kernel_times")
        return tuple(float(v) for v in buf)

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


class DeepOcSortEngine(OcSortEngine):
    """S independent DeepOCSORT trackers (boxmot/trackers/deepocsort/deepocsort.py:265-498) per
    step: the OC-SORT engine plus the embedding term of the first association round (tensor-core
    contraction + adaptive top-2 weighting), the embedding EMA and the CMC correction.

    ReID and CMC are side inputs: ``embs`` [S, det_cap, emb_dim] float32 row-aligned with
    ``dets`` (required unless ``embedding_off``), ``warps`` [S, 2, 3] float64 or None."""

    def __init__(self, n_streams, track_cap=256, det_cap=128, device="cuda:0", det_thresh=0.3, max_age=30,
                 min_hits=3, iou_threshold=0.3, delta_t=3, asso_func="iou", inertia=0.2,
                 w_association_emb=0.5, alpha_fixed_emb=0.95, aw_param=0.5, embedding_off=False,
                 cmc_off=False, aw_off=False, Q_xy_scaling=0.01, Q_s_scaling=0.0001, emb_dim=512,
                 img_wh=(1920, 1080)):
        if not torch.cuda.is_available():
            raise RuntimeError("DeepOcSortEngine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.S, self.Tcap, self.Dcap = int(n_streams), int(track_cap), int(det_cap)
        self.emb_on, self.cmc_off = not embedding_off, bool(cmc_off)
        self.E = int(emb_dim) if self.emb_on else 0
        oc = _lib.OcsortConfig(det_thresh, det_thresh, iou_threshold, inertia, Q_xy_scaling, Q_s_scaling,
                               int(max_age), int(min_hits), int(delta_t), _lib.ASSO_FUNCS[asso_func], 0,
                               int(img_wh[0]), int(img_wh[1]))
        cfg = _lib.DeepocsortConfig(oc, w_association_emb, alpha_fixed_emb, aw_param,
                                    int(bool(embedding_off)), int(bool(aw_off)), self.E)
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_deepocsort_create(ctypes.byref(cfg), self.S, self.Tcap, self.Dcap,
                                                       ctypes.byref(self._h)), "bm_deepocsort_create")
            torch.cuda.synchronize(self.device)
        self.out_rows = min(self.Tcap, self.Dcap)
        self.out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
        self.nout = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
        self.launches = 0

    def step(self, dets, ndets, embs=None, warps=None, out=None, nout=None):
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        if self.emb_on:
            assert embs is not None and embs.dtype == torch.float32 and embs.is_cuda and embs.is_contiguous()
            assert tuple(embs.shape) == (self.S, self.Dcap, self.E)
        if self.cmc_off:
            warps = None
        if warps is not None:
            assert warps.dtype == torch.float64 and warps.is_cuda and tuple(warps.shape) == (self.S, 2, 3)
            assert warps.is_contiguous()
        out = self.out if out is None else out
        nout = self.nout if nout is None else nout
        assert out.dtype == torch.float64 and tuple(out.shape) == (self.S, self.out_rows, 8) and out.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_deepocsort_step(self._h, _ptr(dets), _ptr(ndets),
                                                     _ptr(embs) if self.emb_on else None,
                                                     _ptr(warps) if warps is not None else None,
                                                     _ptr(out), _ptr(nout), _stream_ptr(self.device)),
                       "bm_deepocsort_step")
        self.launches += 6 if self.emb_on else 4
        return out, nout

    def kernel_times(self):
        """(predict+CMC, contraction, assoc, update, finalize, embedding EMA) ms of the last timed step."""
        buf = (ctypes.c_float * 6)()
        _lib.check(_lib.lib().bm_deepocsort_kernel_times(self._h, ctypes.byref(buf)), "bm_deepocsort_kernel_times")
        return tuple(float(v) for v in buf)

    def step_host(self, dets_per_stream, embs_per_stream=None, warps=None):
        """Lists of S arrays (N_s, 6) and (N_s, E) (+ S 2x3 warps); returns a list of (M_s, 8) arrays."""
        if not hasattr(self, "_h_dets"):
            self._h_dets = torch.zeros((self.S, self.Dcap, 6), dtype=torch.float32).pin_memory()
            self._h_nd = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
            self._h_embs = torch.zeros((self.S, self.Dcap, max(self.E, 1)), dtype=torch.float32).pin_memory()
            self._h_out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64).pin_memory()
            self._h_nout = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
        hd, hn, he = self._h_dets.numpy(), self._h_nd.numpy(), self._h_embs.numpy()
        for s, d in enumerate(dets_per_stream):
            n = 0 if d is None else len(d)
            if n > self.Dcap:
                raise RuntimeError("stream %d: %d detections exceed det_cap=%d" % (s, n, self.Dcap))
            if n:
                hd[s, :n] = d
                if self.emb_on:
                    he[s, :n] = embs_per_stream[s]
            hn[s] = n
        d_dets = self._h_dets.to(self.device, non_blocking=True)
        d_nd = self._h_nd.to(self.device, non_blocking=True)
        d_embs = self._h_embs.to(self.device, non_blocking=True) if self.emb_on else None
        d_w = None if warps is None else torch.as_tensor(np.asarray(warps, dtype=np.float64)).reshape(
            self.S, 2, 3).contiguous().to(self.device)
        out, nout = self.step(d_dets, d_nd, d_embs, d_w)
        self._h_out.copy_(out, non_blocking=True)
        self._h_nout.copy_(nout, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        ho, hn2 = self._h_out.numpy(), self._h_nout.numpy()
        return [ho[s, : hn2[s]].copy() for s in range(self.S)]

class BoostTrackEngine:
    """S independent BoostTrack(++) trackers (boxmot/trackers/boosttrack/boosttrack.py:154-336) per
    step: 8-state Kalman filter, DLO / DUO confidence boosts, soft-BIoU / Mahalanobis / shape /
    embedding score matrix, exact assignment, list bookkeeping -- all on the device.

    ReID and CMC are side inputs: ``embs`` [S, det_cap, emb_dim] float32 row-aligned with ``dets``
    (required when ``with_reid``), ``warps`` [S, 2, 3] float64 or None (no camera_update)."""

    def __init__(self, n_streams, track_cap=256, det_cap=128, device="cuda:0", max_age=60, min_hits=3,
                 det_thresh=0.6, iou_threshold=0.3, use_ecc=True, min_box_area=10, aspect_ratio_thresh=1.6,
                 cmc_method="ecc", lambda_iou=0.5, lambda_mhd=0.25, lambda_shape=0.25, use_dlo_boost=True,
                 use_duo_boost=True, dlo_boost_coef=0.65, s_sim_corr=False, use_rich_s=False, use_sb=False,
                 use_vt=False, with_reid=False, emb_dim=512):
        if not torch.cuda.is_available():
            raise RuntimeError("BoostTrackEngine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.S, self.Tcap, self.Dcap = int(n_streams), int(track_cap), int(det_cap)
        self.with_reid, self.use_ecc = bool(with_reid), bool(use_ecc)
        self.E = int(emb_dim) if self.with_reid else 0
        cfg = _lib.BoosttrackConfig(det_thresh, iou_threshold, float(min_box_area), float(aspect_ratio_thresh),
                                    lambda_iou, lambda_mhd, lambda_shape, dlo_boost_coef, int(max_age),
                                    int(min_hits), int(bool(use_dlo_boost)), int(bool(use_duo_boost)),
                                    int(bool(s_sim_corr)), int(bool(use_rich_s)), int(bool(use_sb)),
                                    int(bool(use_vt)), int(self.with_reid), self.E)
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_boosttrack_create(ctypes.byref(cfg), self.S, self.Tcap, self.Dcap,
                                                       ctypes.byref(self._h)), "bm_boosttrack_create")
            torch.cuda.synchronize(self.device)
        self.out_rows = min(self.Tcap, self.Dcap)
        self.out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
        self.nout = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
        self.launches = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            _lib.lib().bm_boosttrack_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_boosttrack_reset(self._h, _stream_ptr(self.device)), "bm_boosttrack_reset")

    def step(self, dets, ndets, embs=None, warps=None, out=None, nout=None):
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        if self.with_reid:
            assert embs is not None and embs.dtype == torch.float32 and embs.is_cuda and embs.is_contiguous()
            assert tuple(embs.shape) == (self.S, self.Dcap, self.E)
        if not self.use_ecc:
            warps = None
        if warps is not None:
            assert warps.dtype == torch.float64 and warps.is_cuda and tuple(warps.shape) == (self.S, 2, 3)
            assert warps.is_contiguous()
        out = self.out if out is None else out
        nout = self.nout if nout is None else nout
        assert out.dtype == torch.float64 and tuple(out.shape) == (self.S, self.out_rows, 8) and out.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_boosttrack_step(self._h, _ptr(dets), _ptr(ndets),
                                                     _ptr(embs) if self.with_reid else None,
                                                     _ptr(warps) if warps is not None else None,
                                                     _ptr(out), _ptr(nout), _stream_ptr(self.device)),
                       "bm_boosttrack_step")
        self.launches += 7 if self.with_reid else 5
        return out, nout

    def select_class(self, n_classes, cls, new_frame):
        """per_class mode (bm_boosttrack_hold_frame): the reference's decorator does not swap BoostTrack's track
        pool, so a class call only rewinds the frame counter.  Asynchronous."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_boosttrack_hold_frame(self._h, int(bool(new_frame)), _stream_ptr(self.device)),
                       "bm_boosttrack_hold_frame")

    def step_classes_host(self, class_dets, class_embs=None, class_warps=None):
        return _step_classes_host(self, class_dets, class_embs, class_warps)

    def counters(self, raise_on_error=True):
        err = ctypes.c_int(0)
        buf = (ctypes.c_ulonglong * 4)()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_boosttrack_counters(self._h, ctypes.byref(err), ctypes.byref(buf),
                                                         _stream_ptr(self.device)), "bm_boosttrack_counters")
        if raise_on_error and err.value:
            raise RuntimeError("BoostTrack engine error: " + "; ".join(m for b, m in ERR_BITS.items() if err.value & b))
        out = dict(zip(("track_updates", "lap_solves", "matches", "births"), [int(v) for v in buf]))
        out["err"] = err.value
        return out

    def status(self, raise_on_error=True):
        return self.counters(raise_on_error)

    def set_stream_groups(self, groups):
        """See OcSortEngine.set_stream_groups."""
        with torch.cuda.device(self.device):
            n = _lib.lib().bm_boosttrack_set_groups(self._h, int(groups))
        if n < 0:
            _lib.check(n, "bm_boosttrack_set_groups")
        self.groups = n
        return n

    def time_kernels(self, enable=True):
        _lib.check(_lib.lib().bm_boosttrack_time_kernels(self._h, int(bool(enable))), "bm_boosttrack_time_kernels")

    def kernel_times(self):
        """(predict, boosts, contraction, assoc+update, finalize, embedding EMA) ms of the last timed step."""
        buf = (ctypes.c_float * 6)()
        _lib.check(_lib.lib().bm_boosttrack_kernel_times(self._h, ctypes.byref(buf)), "bm_boosttrack_kernel_times")
        return tuple(float(v) for v in buf)

    def peek(self, stream_index):
        n = ctypes.c_int(0)
        x = np.zeros((self.Tcap, 8))
        P = np.zeros((self.Tcap, 64))
        ids = np.zeros((self.Tcap,), dtype=np.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_boosttrack_peek(self._h, int(stream_index), ctypes.byref(n), x.ctypes.data,
                                                     P.ctypes.data, ids.ctypes.data, _stream_ptr(self.device)),
                       "bm_boosttrack_peek")
        k = n.value
        return x[:k], P[:k].reshape(k, 8, 8), ids[:k]

    step_host = DeepOcSortEngine.step_host
    emb_on = property(lambda self: self.with_reid)


class HostPipeline:
    """Host <-> device pipeline around an engine's ``step`` for callers with HOST buffers: while frame f is
    being associated, the inputs of frame f+1 (detections, and embeddings / warps for the engines that take
    them) upload on a copy stream from pinned staging buffers; the rows a frame emits are packed on the
    device (``bm_pack_rows``) and written by that kernel STRAIGHT INTO mapped pinned host memory, so the
    device-to-host traffic is 64 bytes per row that exists plus the offsets, not the padded output array.
    ``submit`` returns the finished result of the PREVIOUS frame as ``(rows, offsets)``: stream s owns
    ``rows[offsets[s]:offsets[s + 1]]`` (numpy views of pinned buffers, valid until the next ``submit``)."""

    def __init__(self, engine):
        e = self.eng = engine
        dev = e.device
        self.E = int(getattr(e, "E", 0) or 0) if getattr(e, "emb_on", False) or getattr(e, "with_reid", False) else 0
        self.s_in = torch.cuda.Stream(dev)
        mk = lambda shape, dt: [torch.zeros(shape, dtype=dt, device=dev) for _ in range(2)]
        pin = lambda shape, dt: [torch.zeros(shape, dtype=dt).pin_memory() for _ in range(2)]
        self.d_dets, self.d_nd = mk((e.S, e.Dcap, 6), torch.float32), mk((e.S,), torch.int32)
        self.d_embs = mk((e.S, e.Dcap, self.E), torch.float32) if self.E else None
        self.d_warps = mk((e.S, 2, 3), torch.float64)
        self.d_out, self.d_nout = mk((e.S, e.out_rows, 8), torch.float64), mk((e.S,), torch.int32)
        self.h_rows, self.h_off = pin((e.S * e.out_rows, 8), torch.float64), pin((e.S + 1,), torch.int32)
        self.ev_in = [torch.cuda.Event() for _ in range(2)]
        self.ev_comp = [torch.cuda.Event() for _ in range(2)]
        self.n = 0
        self.h2d_bytes = e.S * e.Dcap * 6 * 4 + e.S * 4 + e.S * e.Dcap * self.E * 4
        self.d2h_rows = 0            # rows written to host memory so far
        self.d2h_frames = 0

    @property
    def d2h_bytes(self):
        """Average bytes the device wrote to host memory per frame: 64 per emitted row + the offsets."""
        f = max(1, self.d2h_frames)
        return self.d2h_rows * 64 // f + (self.eng.S + 1) * 4

    def submit(self, h_dets, h_nd, h_embs=None, h_warps=None):
        """Pinned host tensors of the next frame: h_dets [S,det_cap,6] float32, h_nd [S] int32,
        h_embs [S,det_cap,E] float32 (engines with embeddings), h_warps [S,2,3] float64 or None."""
        b = self.n & 1
        e = self.eng
        cur = torch.cuda.current_stream(e.device)
        if self.n >= 2:
            self.s_in.wait_event(self.ev_comp[b])      # frame n-2 has consumed the device inputs [b]
        with torch.cuda.stream(self.s_in):
            self.d_dets[b].copy_(h_dets, non_blocking=True)
            self.d_nd[b].copy_(h_nd, non_blocking=True)
            if self.E:
                self.d_embs[b].copy_(h_embs, non_blocking=True)
            if h_warps is not None:
                self.d_warps[b].copy_(h_warps, non_blocking=True)
            self.ev_in[b].record(self.s_in)
        cur.wait_event(self.ev_in[b])
        if self.E or h_warps is not None:
            e.step(self.d_dets[b], self.d_nd[b], self.d_embs[b] if self.E else None,
                   self.d_warps[b] if h_warps is not None else None, out=self.d_out[b], nout=self.d_nout[b])
        else:
            e.step(self.d_dets[b], self.d_nd[b], out=self.d_out[b], nout=self.d_nout[b])
        with torch.cuda.device(e.device):
            _lib.check(_lib.lib().bm_pack_rows(_ptr(self.d_out[b]), _ptr(self.d_nout[b]), e.S, e.out_rows,
                                               ctypes.c_void_p(self.h_rows[b].data_ptr()),
                                               ctypes.c_void_p(self.h_off[b].data_ptr()), _stream_ptr(e.device)),
                       "bm_pack_rows")
        self.ev_comp[b].record(cur)
        self.n += 1
        return self._result(1 - b) if self.n >= 2 else None

    def _result(self, b):
        self.ev_comp[b].synchronize()
        off = self.h_off[b].numpy()
        return self.h_rows[b].numpy()[: off[-1]], off

    def drain(self):
        """Result of the last submitted frame."""
        return self._result((self.n - 1) & 1) if self.n else None

    def account(self, result):
        """Bookkeeping for ``d2h_bytes``: call with every result consumed."""
        self.d2h_rows += int(result[1][-1])
        self.d2h_frames += 1


class ByteTrackEngine(_HostStepMixin):
    """S independent ByteTrack trackers advanced by ONE kernel launch per frame.

    Parameters are those of the reference ``ByteTrack.__init__``
    (boxmot/trackers/bytetrack/bytetrack.py:132-156).  Same buffer interface as
    ``OcSortEngine`` (``step`` / ``step_host`` / ``HostPipeline``)."""

    def __init__(self, n_streams, track_cap=256, det_cap=128, device="cuda:0", min_conf=0.1,
                 track_thresh=0.45, match_thresh=0.8, track_buffer=25, frame_rate=30):
        if not torch.cuda.is_available():
            raise RuntimeError("ByteTrackEngine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.S, self.Tcap, self.Dcap = int(n_streams), int(track_cap), int(det_cap)
        cfg = _lib.BytetrackConfig(min_conf, track_thresh, match_thresh, int(track_buffer), int(frame_rate))
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_bytetrack_create(ctypes.byref(cfg), self.S, self.Tcap, self.Dcap,
                                                      ctypes.byref(self._h)), "bm_bytetrack_create")
            torch.cuda.synchronize(self.device)
        self.out_rows = min(self.Tcap, self.Dcap)
        self.out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
        self.nout = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
        self.launches = 0

    def close(self):     # bm_botsort_destroy is the same entry (shared engine state)
        if getattr(self, "_h", None) is not None and self._h.value:
            _lib.lib().bm_bytetrack_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_bytetrack_reset(self._h, _stream_ptr(self.device)), "bm_bytetrack_reset")

    def step(self, dets, ndets, out=None, nout=None):
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        out = self.out if out is None else out
        nout = self.nout if nout is None else nout
        assert out.dtype == torch.float64 and tuple(out.shape) == (self.S, self.out_rows, 8) and out.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_bytetrack_step(self._h, _ptr(dets), _ptr(ndets), _ptr(out), _ptr(nout),
                                                    _stream_ptr(self.device)), "bm_bytetrack_step")
        self.launches += 1
        return out, nout

    def select_class(self, n_classes, cls, new_frame):
        """per_class mode (bm_bytetrack_select_class): make class ``cls``'s active list the engine's; lost / removed
        tracks stay shared, as in the reference's decorator.  Asynchronous."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_bytetrack_select_class(self._h, int(n_classes), int(cls), int(bool(new_frame)),
                                                            _stream_ptr(self.device)), "bm_bytetrack_select_class")

    def step_classes_host(self, class_dets, class_embs=None, class_warps=None):
        return _step_classes_host(self, class_dets, class_embs, class_warps)

    def counters(self, raise_on_error=True):
        """Cumulative {track_updates, lap_solves, matches, births}; raises on device error bits."""
        buf = (ctypes.c_ulonglong * 4)()
        err = ctypes.c_int(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_bytetrack_counters(self._h, ctypes.byref(err), ctypes.byref(buf),
                                                        _stream_ptr(self.device)), "bm_bytetrack_counters")
        if raise_on_error and err.value:
            raise RuntimeError("ByteTrack engine error: " + "; ".join(m for b, m in ERR_BITS.items() if err.value & b))
        d = dict(zip(("track_updates", "lap_solves", "matches", "births"), [int(v) for v in buf]))
        d["err"] = err.value
        return d

    def status(self, raise_on_error=True):
        return self.counters(raise_on_error)

    def peek(self, stream_index):
        """(means [n,8], covariances [n,8,8], ids [n], n_active) -- active then lost tracks."""
        na, nl = ctypes.c_int(0), ctypes.c_int(0)
        m = np.zeros((self.Tcap, 8))
        P = np.zeros((self.Tcap, 64))
        ids = np.zeros((self.Tcap,), dtype=np.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_bytetrack_peek(self._h, int(stream_index), ctypes.byref(na), ctypes.byref(nl),
                                                    m.ctypes.data, P.ctypes.data, ids.ctypes.data,
                                                    _stream_ptr(self.device)), "bm_bytetrack_peek")
        k = na.value + nl.value
        return m[:k], P[:k].reshape(k, 8, 8), ids[:k], na.value


class BotSortEngine(ByteTrackEngine):
    """S independent BoT-SORT trackers (boxmot/trackers/botsort/botsort.py:49-411) per step.

    ReID and CMC are side inputs: ``embs`` [S, det_cap, emb_dim] float32 row-aligned with
    ``dets`` (required when ``with_reid``), ``warps`` [S, 2, 3] float64 or None (identity)."""

    def __init__(self, n_streams, track_cap=256, det_cap=128, device="cuda:0", track_high_thresh=0.5,
                 track_low_thresh=0.1, new_track_thresh=0.6, track_buffer=30, match_thresh=0.8,
                 proximity_thresh=0.5, appearance_thresh=0.25, frame_rate=30, fuse_first_associate=False,
                 with_reid=True, emb_dim=512):
        if not torch.cuda.is_available():
            raise RuntimeError("BotSortEngine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.S, self.Tcap, self.Dcap = int(n_streams), int(track_cap), int(det_cap)
        self.with_reid, self.E = bool(with_reid), int(emb_dim) if with_reid else 0
        cfg = _lib.BotsortConfig(track_high_thresh, track_low_thresh, new_track_thresh, match_thresh,
                                 proximity_thresh, appearance_thresh, int(track_buffer), int(frame_rate),
                                 int(bool(fuse_first_associate)), int(self.with_reid), self.E)
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_botsort_create(ctypes.byref(cfg), self.S, self.Tcap, self.Dcap,
                                                    ctypes.byref(self._h)), "bm_botsort_create")
            torch.cuda.synchronize(self.device)
        self.out_rows = min(self.Tcap, self.Dcap)
        self.out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
        self.nout = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
        self.launches = 0

    def set_stream_groups(self, groups):
        """See OcSortEngine.set_stream_groups."""
        with torch.cuda.device(self.device):
            n = _lib.lib().bm_botsort_set_groups(self._h, int(groups))
        if n < 0:
            _lib.check(n, "bm_botsort_set_groups")
        self.groups = n
        return n

    def step(self, dets, ndets, embs=None, warps=None, out=None, nout=None):
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        if self.with_reid:
            assert embs is not None and embs.dtype == torch.float32 and embs.is_cuda and embs.is_contiguous()
            assert tuple(embs.shape) == (self.S, self.Dcap, self.E)
        if warps is not None:
            assert warps.dtype == torch.float64 and warps.is_cuda and tuple(warps.shape) == (self.S, 2, 3)
            assert warps.is_contiguous()
        out = self.out if out is None else out
        nout = self.nout if nout is None else nout
        assert out.dtype == torch.float64 and tuple(out.shape) == (self.S, self.out_rows, 8) and out.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_botsort_step(self._h, _ptr(dets), _ptr(ndets),
                                                  _ptr(embs) if self.with_reid else None,
                                                  _ptr(warps) if warps is not None else None,
                                                  _ptr(out), _ptr(nout), _stream_ptr(self.device)), "bm_botsort_step")
        self.launches += 3 if self.with_reid else 1
        return out, nout

    def step_host(self, dets_per_stream, embs_per_stream=None, warps=None):
        """Lists of S arrays (N_s, 6) and (N_s, E); returns a list of (M_s, 8) arrays."""
        if not hasattr(self, "_h_dets"):
            self._h_dets = torch.zeros((self.S, self.Dcap, 6), dtype=torch.float32).pin_memory()
            self._h_nd = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
            self._h_embs = torch.zeros((self.S, self.Dcap, max(self.E, 1)), dtype=torch.float32).pin_memory()
            self._h_out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64).pin_memory()
            self._h_nout = torch.zeros((self.S,), dtype=torch.int32).pin_memory()
        hd, hn, he = self._h_dets.numpy(), self._h_nd.numpy(), self._h_embs.numpy()
        for s, d in enumerate(dets_per_stream):
            n = 0 if d is None else len(d)
            if n > self.Dcap:
                raise RuntimeError("stream %d: %d detections exceed det_cap=%d" % (s, n, self.Dcap))
            if n:
                hd[s, :n] = d
                if self.with_reid:
                    he[s, :n] = embs_per_stream[s]
            hn[s] = n
        d_dets = self._h_dets.to(self.device, non_blocking=True)
        d_nd = self._h_nd.to(self.device, non_blocking=True)
        d_embs = self._h_embs.to(self.device, non_blocking=True) if self.with_reid else None
        d_w = None if warps is None else torch.as_tensor(np.asarray(warps, dtype=np.float64)).reshape(
            self.S, 2, 3).contiguous().to(self.device)
        out, nout = self.step(d_dets, d_nd, d_embs, d_w)
        self._h_out.copy_(out, non_blocking=True)
        self._h_nout.copy_(nout, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        ho, hn2 = self._h_out.numpy(), self._h_nout.numpy()
        return [ho[s, : hn2[s]].copy() for s in range(self.S)]


SS_ERR_BITS = {1: "track capacity exceeded", 2: "detection capacity exceeded", 8: "feature pool capacity exceeded",
               16: "non-finite costs reached the assignment solver"}


class StrongSortEngine:
    """S independent StrongSORT trackers (the fork's "enhanced" StrongSORT, boxmot/trackers/strongsort/
    strongsort.py:115-205 over sort/tracker.py, sort/track.py, sort/linear_assignment.py) per step, all on
    the device: detection quality sort, camera_update + Kalman predict, the sample-bank cosine metric on the
    tensor cores, Mahalanobis-gated three-stage cascade with scipy-exact assignments, adaptive feature EMA,
    id recovery from the lost buffer, partial_fit.

    The reference runs only in one configuration: occlusion handling and crowd detection off,
    and tracks born Confirmed as in its own CI mode (``born_confirmed=True``).  ReID and CMC are side inputs:
    ``embs`` [S, det_cap, emb_dim] float32 (required), ``warps`` [S, 2, 3] float64 or None (identity)."""

    def __init__(self, n_streams, track_cap=512, det_cap=128, pool_cap=None, device="cuda:0", min_conf=0.1,
                 max_cos_dist=0.15, max_iou_dist=0.7, max_age=50, n_init=2, nn_budget=150, mc_lambda=0.995,
                 ema_alpha=0.9, conf_thresh_high=0.7, conf_thresh_low=0.3, id_preservation_weight=0.1,
                 born_confirmed=True, emb_dim=512, **_ignored):
        if not torch.cuda.is_available():
            raise RuntimeError("StrongSortEngine needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        self.S, self.Tcap, self.Dcap, self.E = int(n_streams), int(track_cap), int(det_cap), int(emb_dim)
        self.Pcap = int(pool_cap) if pool_cap else 4 * self.Tcap
        cfg = _lib.StrongsortConfig(min_conf, max_cos_dist, max_iou_dist, mc_lambda, ema_alpha, conf_thresh_high,
                                    conf_thresh_low, id_preservation_weight, int(max_age), int(n_init),
                                    int(nn_budget), int(bool(born_confirmed)), self.E)
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_strongsort_create(ctypes.byref(cfg), self.S, self.Tcap, self.Dcap, self.Pcap,
                                                       ctypes.byref(self._h)), "bm_strongsort_create")
        self.out_rows = self.Dcap
        self.out = torch.zeros((self.S, self.out_rows, 8), dtype=torch.float64, device=self.device)
        self.nout = torch.zeros((self.S,), dtype=torch.int32, device=self.device)
        self.launches = 0
        self.with_reid = True

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            _lib.lib().bm_strongsort_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_strongsort_reset(self._h, _stream_ptr(self.device)), "bm_strongsort_reset")

    def step(self, dets, ndets, embs=None, warps=None, out=None, nout=None):
        assert dets.dtype == torch.float32 and dets.is_cuda and dets.is_contiguous()
        assert tuple(dets.shape) == (self.S, self.Dcap, 6)
        assert ndets.dtype == torch.int32 and ndets.is_cuda and ndets.numel() == self.S
        assert embs is not None and embs.dtype == torch.float32 and embs.is_cuda and embs.is_contiguous()
        assert tuple(embs.shape) == (self.S, self.Dcap, self.E)
        if warps is not None:
            assert warps.dtype == torch.float64 and warps.is_cuda and tuple(warps.shape) == (self.S, 2, 3)
            assert warps.is_contiguous()
        out = self.out if out is None else out
        nout = self.nout if nout is None else nout
        assert out.dtype == torch.float64 and tuple(out.shape) == (self.S, self.out_rows, 8) and out.is_contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_strongsort_step(self._h, _ptr(dets), _ptr(ndets), _ptr(embs),
                                                     _ptr(warps) if warps is not None else None,
                                                     _ptr(out), _ptr(nout), _stream_ptr(self.device)),
                       "bm_strongsort_step")
        self.launches += 6
        return out, nout

    def counters(self, raise_on_error=True):
        err = ctypes.c_int(0)
        buf = (ctypes.c_ulonglong * 4)()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_strongsort_counters(self._h, ctypes.byref(err), ctypes.byref(buf),
                                                         _stream_ptr(self.device)), "bm_strongsort_counters")
        if raise_on_error and err.value:
            raise RuntimeError("StrongSORT engine error: " + "; ".join(m for b, m in SS_ERR_BITS.items() if err.value & b))
        out = dict(zip(("track_updates", "lap_solves", "matches", "births"), [int(v) for v in buf]))
        out["err"] = err.value
        return out

    def status(self, raise_on_error=True):
        return self.counters(raise_on_error)

    def set_stream_groups(self, groups):
        """Step the streams as ``groups`` independent groups on CUDA streams of their own (forked from / joined
        into the current stream inside ``step``); returns the number in use.  Results do not depend on it."""
        with torch.cuda.device(self.device):
            n = _lib.lib().bm_strongsort_set_groups(self._h, int(groups))
        if n < 0:
            _lib.check(n, "bm_strongsort_set_groups")
        self.groups = n
        return n

    def usage(self):
        """Over all streams: (track slots in use, pool slots in use, vectors referenced by a sample list)."""
        a, b, c = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_strongsort_usage(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c),
                                                      _stream_ptr(self.device)), "bm_strongsort_usage")
        return {"max_tracks": a.value, "max_pool_used": b.value, "max_live": c.value}

    def time_kernels(self, enable=True):
        _lib.check(_lib.lib().bm_strongsort_time_kernels(self._h, int(bool(enable))), "bm_strongsort_time_kernels")

    def kernel_times(self):
        """(prepare+predict, bank contraction, cost matrix, cascade, updates, finalize) ms of the last timed step."""
        buf = (ctypes.c_float * 6)()
        _lib.check(_lib.lib().bm_strongsort_kernel_times(self._h, ctypes.byref(buf)), "bm_strongsort_kernel_times")
        return tuple(float(v) for v in buf)

    def peek(self, stream_index):
        """(means [n,8], covariances [n,8,8], table rows [n,42], ids [n], sample-list lengths [n], n_lost)."""
        n, nl = ctypes.c_int(0), ctypes.c_int(0)
        x = np.zeros((self.Tcap, 8))
        P = np.zeros((self.Tcap, 64))
        tab = np.zeros((self.Tcap, 42))
        ids = np.zeros((self.Tcap,), dtype=np.int32)
        nb = np.zeros((self.Tcap,), dtype=np.int32)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bm_strongsort_peek(self._h, int(stream_index), ctypes.byref(n), ctypes.byref(nl),
                                                     x.ctypes.data, P.ctypes.data, tab.ctypes.data, ids.ctypes.data,
                                                     nb.ctypes.data, _stream_ptr(self.device)), "bm_strongsort_peek")
        k = n.value
        return x[:k], P[:k].reshape(k, 8, 8), tab[:k], ids[:k], nb[:k], nl.value

    step_host = DeepOcSortEngine.step_host
    emb_on = property(lambda self: True)
