"""ByteTrack behind the reference plugin API (boxmot/trackers/bytetrack/bytetrack.py:119-302).

``ByteTrack`` is the single-stream adapter; ``boxmot_b200.engine.ByteTrackEngine`` is the
throughput path that advances thousands of streams per launch."""
import numpy as np

from boxmot_b200.engine import ByteTrackEngine
from boxmot_b200.trackers.basetracker import BaseTracker


class ByteTrack(BaseTracker):
    def __init__(self, min_conf=0.1, track_thresh=0.45, match_thresh=0.8, track_buffer=25, frame_rate=30,
                 per_class=False, track_cap=512, det_cap=512, device="cuda:0"):
        super().__init__(per_class=bool(per_class))
        self.lost_stracks = []
        self.removed_stracks = []
        self.frame_id = 0
        self.track_buffer = track_buffer
        self.min_conf = min_conf
        self.track_thresh = track_thresh
        self.match_thresh = match_thresh
        self.det_thresh = track_thresh
        self.buffer_size = int(frame_rate / 30.0 * track_buffer)
        self.max_time_lost = self.buffer_size
        self._frame_rate = frame_rate
        self._caps = (track_cap, det_cap)
        self._device = device
        self._engine = None

    def _make_engine(self):
        self._engine = ByteTrackEngine(
            1, track_cap=self._caps[0], det_cap=self._caps[1], device=self._device,
            min_conf=self.min_conf, track_thresh=self.track_thresh, match_thresh=self.match_thresh,
            track_buffer=self.track_buffer, frame_rate=self._frame_rate)

    def update(self, dets, img=None, embs=None):
        dets = self._prepare(dets, img)
        self.check_inputs(dets, img)
        if self._engine is None:
            self._make_engine()
        self.frame_count += 1
        if self.per_class:
            # per_class_decorator (basetracker.py:154-201): update once per class id with that class's active_tracks;
            # lost_stracks / removed_stracks are ONE list for all classes (bytetrack.py:143-144), so every class call
            # also predicts, re-finds and ages the lost tracks of the others -- reproduced on the device
            # (bm_bytetrack_select_class).  det_ind is the index within the class subset.
            per_cls = [dets[dets[:, 5] == c] if dets.size else dets[:0] for c in range(self.nr_classes)]
            rows = [r for r in self._engine.step_classes_host(per_cls) if len(r)]
            self._engine.status()
            return np.vstack(rows) if rows else np.empty((0, 8))
        rows = self._engine.step_host([dets])[0]
        self._engine.status()
        # np.asarray(outputs) of an empty list (bytetrack.py:301)
        return rows if len(rows) else np.asarray([])

    def reset(self):
        if self._engine is not None:
            self._engine.reset()
        self.frame_count = 0
