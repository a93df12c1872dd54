"""MOT result rows / files against the reference's own convert_to_mot_format +
write_mot_results (boxmot/engine/utils.py:101-172) run on the ByteTrack MOT17-02-FRCNN golden rows
(tests/golden/ref_mot_results_bytetrack_MOT17_02_FRCNN.txt, oracle/gen_golden.py --mot-writer): byte-identical text."""
import os

import numpy as np
import pytest

from conftest import GOLD

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _golden():
    rows = np.load(os.path.join(GOLD, "ref_bytetrack_MOT17_02_FRCNN.npz"))["rows"]
    want = open(os.path.join(GOLD, "ref_mot_results_bytetrack_MOT17_02_FRCNN.txt")).read()
    return rows, want


def test_convert_and_write_mot_results_match_the_reference_text(tmp_path):
    from boxmot_b200.mot_io import convert_to_mot_format, write_mot_results
    rows, want = _golden()
    path = tmp_path / "out" / "MOT17-02-FRCNN.txt"
    for f in range(int(rows[:, 0].max()) + 1):
        r = rows[rows[:, 0] == f][:, 1:]
        m = convert_to_mot_format(r, f + 1)
        assert (m is None) == (r.size == 0)
        write_mot_results(path, m)
    assert path.read_text() == want


def test_batch_writer_splits_streams_and_matches_the_reference_text(tmp_path):
    """Three streams per step (the golden rows, nothing, the golden rows again) through bm_pack_mot_rows
    and MotBatchWriter: one file per stream, each byte-identical to the reference's."""
    from boxmot_b200.mot_io import MotBatchWriter
    rows, want = _golden()
    paths = [tmp_path / ("s%d.txt" % s) for s in range(3)]
    w = MotBatchWriter(paths)
    R = 64
    for f in range(int(rows[:, 0].max()) + 1):
        r = rows[rows[:, 0] == f][:, 1:]
        out = np.zeros((3, R, 8))
        out[0, : len(r)] = r
        out[2, : len(r)] = r
        nout = np.array([len(r), 0, len(r)], np.int32)
        w.append(torch.as_tensor(out).cuda(), torch.as_tensor(nout).cuda(), f + 1)
    assert paths[0].read_text() == want and paths[2].read_text() == want and paths[1].read_text() == ""


def test_mot_rounding_is_half_to_even_like_numpy():
    from boxmot_b200.mot_io import convert_to_mot_format
    r = np.array([[0.5, 1.5, 3.0, 5.0, 7, 0.25, 2, 0], [2.5, -0.5, 6.0, 3.0, 8, 0.5, 0, 1]], float)
    m = convert_to_mot_format(r, 3)
    want = np.column_stack((np.full((2, 1), 3), r[:, 4].astype(np.int32),
                            np.c_[r[:, 0], r[:, 1], r[:, 2] - r[:, 0], r[:, 3] - r[:, 1]].round().astype(np.int32),
                            np.ones((2, 1), np.int32), r[:, 6].astype(np.int32), r[:, 5]))
    assert np.array_equal(m, want)
