"""MOT-challenge result rows and files for the engines' outputs (the step right after the
association path).  Mirrors boxmot/engine/utils.py:101-172: ``convert_to_mot_format(results, frame_idx)``
(ndarray branch) and ``write_mot_results(txt_path, mot_results)``; the row conversion runs on the device for
all streams of a frame at once (``bm_pack_mot_rows``: xyxy -> ltwh, np.round's half-to-even rounding, int32
casts, ragged packing straight into pinned host memory), the text is the reference's own format string."""
import ctypes
import io
from pathlib import Path

import numpy as np
import torch

from boxmot_b200 import _lib

MOT_FMT = "%d,%d,%d,%d,%d,%d,%d,%d,%.6f"      # boxmot/engine/utils.py:172


def convert_batch_to_mot_format(d_out, d_nout, frame_idx, h_rows=None, h_off=None):
    """d_out [S, out_rows, 8] float64 / d_nout [S] int32 device tensors of one engine step ->
    (mot_rows [sum(nout), 9] float64, offsets [S + 1] int32) as numpy views of pinned host buffers."""
    S, out_rows = int(d_out.shape[0]), int(d_out.shape[1])
    if h_rows is None:
        h_rows = torch.zeros((S * out_rows, 9), dtype=torch.float64).pin_memory()
        h_off = torch.zeros((S + 1,), dtype=torch.int32).pin_memory()
    dev = d_out.device
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev)
        _lib.check(_lib.lib().bm_pack_mot_rows(ctypes.c_void_p(d_out.data_ptr()), ctypes.c_void_p(d_nout.data_ptr()), S,
                                               out_rows, int(frame_idx), ctypes.c_void_p(h_rows.data_ptr()),
                                               ctypes.c_void_p(h_off.data_ptr()), ctypes.c_void_p(st.cuda_stream)),
                   "bm_pack_mot_rows")
        st.synchronize()
    off = h_off.numpy()
    return h_rows.numpy()[: off[-1]], off


def convert_to_mot_format(results, frame_idx):
    """The reference's signature for ONE tracker's (M, 7+) result array of a frame; returns (M, 9) float64
    rows [frame, id, l, t, w, h, 1, cls, conf] (None for an empty result, as the reference falls through)."""
    results = np.asarray(results)
    if results.size == 0:
        return None
    r = np.zeros((1, len(results), 8))
    r[0, :, :7] = results[:, :7]
    d_out = torch.as_tensor(r).cuda()
    d_n = torch.as_tensor(np.array([len(results)], np.int32)).cuda()
    rows, _ = convert_batch_to_mot_format(d_out, d_n, frame_idx)
    return rows.copy()


def write_mot_results(txt_path, mot_results):
    """boxmot/engine/utils.py:149-172: append the rows to the text file (created with its directory)."""
    if mot_results is None:
        return
    txt_path = Path(txt_path)
    txt_path.parent.mkdir(parents=True, exist_ok=True)
    txt_path.touch(exist_ok=True)
    if mot_results.size != 0:
        with open(str(txt_path), "a") as f:
            np.savetxt(f, mot_results, fmt=MOT_FMT)


class MotBatchWriter:
    """One MOT result file per stream, appended frame by frame from the packed rows of a step: the whole
    frame is formatted once and the text is split at the streams' row offsets."""

    def __init__(self, paths):
        self.paths = [Path(p) for p in paths]
        for p in self.paths:
            p.parent.mkdir(parents=True, exist_ok=True)
            p.write_text("")
        S = len(self.paths)
        self._bufs = None
        self.S = S

    def append(self, d_out, d_nout, frame_idx):
        if self._bufs is None:
            self._bufs = (torch.zeros((self.S * int(d_out.shape[1]), 9), dtype=torch.float64).pin_memory(),
                          torch.zeros((self.S + 1,), dtype=torch.int32).pin_memory())
        rows, off = convert_batch_to_mot_format(d_out, d_nout, frame_idx, *self._bufs)
        if not len(rows):
            return 0
        buf = io.StringIO()
        np.savetxt(buf, rows, fmt=MOT_FMT)
        lines = buf.getvalue().splitlines(keepends=True)
        for s, p in enumerate(self.paths):
            if off[s + 1] > off[s]:
                with open(p, "a") as f:
                    f.writelines(lines[off[s]:off[s + 1]])
        return len(rows)
