"""Reference-facing tracker base: the drop-in boundary of the association path.

Mirrors what callers of the reference rely on (boxmot/trackers/basetracker.py:12-81,
:108-226): constructor attributes, ``update(dets, img, embs)`` input handling, the
``AssertionError`` input checks and the float32 round trip every ndarray goes through
(:122-128, ndarray has a ``.data`` memoryview).  Plotting helpers (:228-409) are out of
scope for this package."""
import numpy as np


class BaseTracker:
    def __init__(self, det_thresh=0.3, max_age=30, min_hits=3, iou_threshold=0.3, max_obs=50,
                 nr_classes=80, per_class=False, asso_func="iou", is_obb=False):
        self.det_thresh = det_thresh
        self.max_age = max_age
        self.max_obs = max_obs
        self.min_hits = min_hits
        self.per_class = per_class
        self.nr_classes = nr_classes
        self.iou_threshold = iou_threshold
        self.last_emb_size = None
        self.asso_func_name = asso_func + "_obb" if is_obb else asso_func
        self.is_obb = is_obb
        self.frame_count = 0
        self.active_tracks = []
        self.per_class_active_tracks = None
        self._first_frame_processed = False
        self._first_dets_processed = False
        if self.per_class:
            self.per_class_active_tracks = {i: [] for i in range(self.nr_classes)}
        if self.max_age >= self.max_obs:
            self.max_obs = self.max_age + 5

    # -- input handling shared by every tracker ---------------------------------------
    def _prepare(self, dets, img):
        """setup_decorator + per_class_decorator input normalisation
        (basetracker.py:116-163)."""
        if hasattr(dets, "data") and not isinstance(dets, np.ndarray):
            dets = dets.data
        if isinstance(dets, np.ndarray) and dets.dtype != object and dets.size:
            dets = np.array(dets, dtype=np.float32)   # the reference's memoryview branch
        if not self._first_dets_processed and dets is not None and getattr(dets, "ndim", 0) == 2:
            if dets.shape[1] == 6:
                self.is_obb = False
                self._first_dets_processed = True
            elif dets.shape[1] == 7:
                self.is_obb = True
                self._first_dets_processed = True
        if not self._first_frame_processed and img is not None:
            self.h, self.w = img.shape[0:2]
            self._first_frame_processed = True
        if dets is None or len(dets) == 0:
            dets = np.empty((0, 6))
        return dets

    def check_inputs(self, dets, img, embs=None):
        assert isinstance(dets, np.ndarray), \
            f"Unsupported 'dets' input format '{type(dets)}', valid format is np.ndarray"
        assert isinstance(img, np.ndarray), \
            f"Unsupported 'img_numpy' input format '{type(img)}', valid format is np.ndarray"
        assert len(dets.shape) == 2, \
            "Unsupported 'dets' dimensions, valid number of dimensions is two"
        if embs is not None:
            assert dets.shape[0] == embs.shape[0], \
                "Missmatch between detections and embeddings sizes"
        if self.is_obb:
            assert dets.shape[1] == 7, \
                "Unsupported 'dets' 2nd dimension lenght, valid lenghts is 6 (cx,cy,w,h,angle,conf,cls)"
        else:
            assert dets.shape[1] == 6, \
                "Unsupported 'dets' 2nd dimension lenght, valid lenghts is 6 (x1,y1,x2,y2,conf,cls)"

    def update(self, dets, img, embs=None):
        raise NotImplementedError("The update method needs to be implemented by the subclass.")

    def reset(self):
        pass
