// Device state of the stream-batched ByteTrack engine (one STrack per slot, SoA by slot).
// Mirrors the objects of boxmot/trackers/bytetrack/bytetrack.py: STrack (:14-116) fields and
// the tracker's three lists (active_tracks, lost_stracks, removed_stracks; :143-145).
#pragma once
#include "common.cuh"
#include "kf_xyah.cuh"

namespace bm {

// per-slot integers
enum { BI_ID = 0, BI_STATE, BI_ACTIVATED, BI_FRAME, BI_START, BI_DETIND, BI_INREMOVED, BI_COUNT = 8 };
// TrackState (bytetrack/basetrack.py:8-12)
enum { BT_NEW = 0, BT_TRACKED = 1, BT_LOST = 2, BT_REMOVED = 3 };
// per-stream header
enum { BH_NACT = 0, BH_NLOST, BH_NEXT_ID, BH_FRAME, BH_ERR, BH_LAPS, BH_MATCH, BH_BIRTHS, BH_COUNT = 8 };

struct ByteParams {
    double track_thresh, min_conf, match_thresh, det_thresh;
    int max_time_lost;
};

struct ByteState {
    int S, Tcap, Dcap;
    double* xP;        // [S][XA_N][Tcap]  mean(8) + block covariance(12)
    double* confcls;   // [S][2][Tcap]
    double* tbox;      // [S][4][Tcap]     STrack.xyxy of the slot, refreshed where it is read
    int* ints;         // [S][BI_COUNT][Tcap]
    int* act;          // [S][Tcap] active_tracks, in list order (slots)
    int* lost;         // [S][Tcap] lost_stracks, in list order
    int* hdr;          // [S][BH_COUNT]
    unsigned long long* n_updates;   // [S] pool tracks that entered predict
    double* scratch;   // [S][Tcap*Dcap] materialised cost matrices (L2-resident)
};

size_t bytetrack_step_smem(const ByteState& st);
cudaError_t bytetrack_step_configure(size_t smem);
cudaError_t bytetrack_step_launch(const ByteState& st, const ByteParams& pr, const float* dets, const int* ndets,
                                  double* out, int* nout, cudaStream_t stream);

}  // namespace bm
