"""OC-SORT behind the reference plugin API (boxmot/trackers/ocsort/ocsort.py:195-439).

``OcSort`` is the single-stream adapter (one engine stream per tracker, or one per class
when ``per_class``); ``boxmot_b200.engine.OcSortEngine`` is the throughput path that
advances thousands of streams per launch."""
import numpy as np

from boxmot_b200.engine import OcSortEngine
from boxmot_b200.trackers.basetracker import BaseTracker


class OcSort(BaseTracker):
    def __init__(self, per_class=False, min_conf=0.1, det_thresh=0.2, max_age=30, min_hits=3,
                 asso_threshold=0.3, delta_t=3, asso_func="iou", inertia=0.2, use_byte=False,
                 Q_xy_scaling=0.01, Q_s_scaling=0.0001, track_cap=512, det_cap=512,
                 device="cuda:0"):
        super().__init__(max_age=max_age, per_class=bool(per_class), asso_func=asso_func)
        self.min_conf = min_conf
        self.min_hits = min_hits
        self.asso_threshold = asso_threshold
        self.det_thresh = det_thresh
        self.delta_t = delta_t
        self.inertia = inertia
        self.use_byte = use_byte
        self.Q_xy_scaling = Q_xy_scaling
        self.Q_s_scaling = Q_s_scaling
        self._caps = (track_cap, det_cap)
        self._device = device
        self._engine = None

    def _make_engine(self):
        # per_class (basetracker.py:154-201): one engine stream per class id, same frame count
        n = self.nr_classes if self.per_class else 1
        self._created = np.zeros((n,), dtype=np.int64)      # tracks created so far, per class
        self._global_ids = [[] for _ in range(n)]           # class-local id -> tracker-wide id
        self._count = 0                                      # KalmanBoxTracker.count
        self._engine = OcSortEngine(
            n, track_cap=self._caps[0], det_cap=self._caps[1], device=self._device,
            min_conf=self.min_conf, det_thresh=self.det_thresh, max_age=self.max_age,
            min_hits=self.min_hits, asso_threshold=self.asso_threshold, delta_t=self.delta_t,
            asso_func=self.asso_func_name, inertia=self.inertia, use_byte=self.use_byte,
            Q_xy_scaling=self.Q_xy_scaling, Q_s_scaling=self.Q_s_scaling,
            img_wh=(getattr(self, "w", 1920), getattr(self, "h", 1080)))

    def update(self, dets, img, embs=None):
        dets = self._prepare(dets, img)
        self.check_inputs(dets, img)
        if self._engine is None:
            self._make_engine()
        self.frame_count += 1
        if not self.per_class:
            rows = self._engine.step_host([dets])[0]
            self._engine.status()
            return rows if len(rows) else np.array([])
        # get_class_dets_n_embs (basetracker.py:83-106): rows of class c, in input order;
        # det_ind is then the index WITHIN the class subset (ocsort.py:265 runs per class)
        per_cls = [dets[dets[:, 5] == c] if dets.size else dets[:0] for c in range(self.nr_classes)]
        rows = self._engine.step_host(per_cls)
        self._engine.status()
        # ids: the reference draws every new track's id from ONE counter while it walks the
        # classes in order (KalmanBoxTracker.count, ocsort.py:79-80), emitted or not
        _, created = self._engine.stream_info()
        for c in range(self.nr_classes):
            for _ in range(int(created[c] - self._created[c])):
                self._global_ids[c].append(self._count)
                self._count += 1
        self._created[:] = created
        out = []
        for c, r in enumerate(rows):
            if len(r):
                r = r.copy()
                r[:, 4] = [self._global_ids[c][int(i) - 1] + 1 for i in r[:, 4]]
                out.append(r)
        return np.vstack(out) if out else np.empty((0, 8))

    def reset(self):
        if self._engine is not None:
            self._engine.reset()
            self._created[:] = 0
            self._global_ids = [[] for _ in self._global_ids]
            self._count = 0
        self.frame_count = 0
