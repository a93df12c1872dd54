// Device state of the stream-batched ByteTrack engine (one STrack per slot, SoA by slot).
// Mirrors the objects of boxmot/trackers/bytetrack/bytetrack.py: STrack (:14-116) fields and
// the tracker's three lists (active_tracks, lost_stracks, removed_stracks; :143-145).
#pragma once
#include "common.cuh"
#include "kf_xyah.cuh"

namespace bm {

// per-slot integers
enum { BI_ID = 0, BI_STATE, BI_ACTIVATED, BI_FRAME, BI_START, BI_DETIND, BI_INREMOVED, BI_NHIST,
       BI_DENSE,          // BoT-SORT: the covariance lives in the dense side array dP (after a rotating warp)
       BI_COUNT };
// TrackState (bytetrack/basetrack.py:8-12)
enum { BT_NEW = 0, BT_TRACKED = 1, BT_LOST = 2, BT_REMOVED = 3 };
// per-stream header
enum { BH_NACT = 0, BH_NLOST, BH_NEXT_ID, BH_FRAME, BH_ERR, BH_LAPS, BH_MATCH, BH_BIRTHS, BH_NFUPD, BH_NLIVE, BH_NH,
       BH_COUNT = 12 };
constexpr int BOT_HIST = 4;          // distinct classes remembered per track (botsort_track.py:51-64)
constexpr int BM_ERR_CLS_HIST = 16;  // a track saw more than BOT_HIST distinct classes
constexpr int BM_ERR_WARP = 32;      // camera-motion warp with rotation / shear (needs a dense covariance)

struct ByteParams {
    double track_thresh, min_conf, match_thresh, det_thresh;   // det_thresh: births need conf >= it
    int max_time_lost;
    // BoT-SORT only (botsort.py:49-92)
    double proximity_thresh, appearance_thresh;
    int fuse_first, with_reid, emb_dim;
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
    // BoT-SORT only
    float* feat;       // [S][Tcap][E]  smooth_feat of the slot (botsort_track.py:40-49)
    float* dfeat;      // [S][Dcap][E]  this frame's detection features as STrack.__init__ leaves them
    float* embd;       // [S][Tcap][Dcap] max(0, cosine distance) of (slot, high-score detection position)
    double* chist;     // [S][2*BOT_HIST][Tcap] class votes: class ids, then accumulated confidences
    double* dP;        // [S][64][Tcap] dense 8x8 covariance of the tracks with BI_DENSE (multi_gmc with rotation)
    int* fupd;         // [S][2][Dcap] (slot, detection row | birth flag << 30) feature updates of the frame
    int* elive;        // [S][Tcap] live slots (active then lost list) = rows of the embedding contraction
    int* ehind;        // [S][Dcap] rows of the high-score detections = its columns
    float* etn;        // [S][Tcap] |smooth_feat| by row position
    float* edn;        // [S][Dcap] |detection feature| by column position
    int tensor_cores;  // 1: tcgen05 contraction (csrc/cosine_tc.cuh), 0: fp32 FFMA tiles in bot_embed_kernel
    // per_class mode (bm_bytetrack_select_class): the active lists of the classes that are not the current one; their
    // slots are in use although they are in neither st.act nor st.lost
    int* pc_act = nullptr;    // [S][pc_C][Tcap]
    int* pc_n = nullptr;      // [S][pc_C]
    int* pc_meta = nullptr;   // [S][2] current class (-1: none), frame counter at the frame's start
    int pc_C = 0;
};

// Streams [s0, s0 + n) as an engine of their own (every array is stream-major); see ocs_substate.
inline ByteState byte_substate(const ByteState& st, int s0, int n, int E) {
    ByteState g = st;
    const size_t o = (size_t)s0, T = st.Tcap, D = st.Dcap;
    g.S = n;
    if (g.pc_act) { g.pc_act += o * st.pc_C * T; g.pc_n += o * st.pc_C; g.pc_meta += o * 2; }
    g.xP += o * XA_N * T;  g.confcls += o * 2 * T;  g.tbox += o * 4 * T;  g.ints += o * BI_COUNT * T;
    g.act += o * T;  g.lost += o * T;  g.hdr += o * BH_COUNT;  g.n_updates += o;  g.scratch += o * T * D;
    if (st.chist) g.chist += o * 2 * BOT_HIST * T;
    if (st.dP) g.dP += o * 64 * T;
    if (st.fupd) g.fupd += o * 2 * D;
    if (st.feat) g.feat += o * T * E;
    if (st.dfeat) g.dfeat += o * D * E;
    if (st.embd) g.embd += o * T * D;
    if (st.elive) g.elive += o * T;
    if (st.ehind) g.ehind += o * D;
    if (st.etn) g.etn += o * T;
    if (st.edn) g.edn += o * D;
    return g;
}

size_t bytetrack_step_smem(const ByteState& st);
cudaError_t bytetrack_step_configure(size_t smem);
cudaError_t bytetrack_step_launch(const ByteState& st, const ByteParams& pr, const float* dets, const int* ndets,
                                  double* out, int* nout, cudaStream_t stream);
// BoT-SORT: embedding kernel (split + feature preparation + cosine GEMM), the step, feature EMA.
// embs [S][Dcap][E] float32 (may be NULL when with_reid = 0); warps [S][6] double or NULL (identity).
cudaError_t botsort_step_launch(const ByteState& st, const ByteParams& pr, const float* dets, const int* ndets,
                                const float* embs, const double* warps, double* out, int* nout, cudaStream_t stream);

}  // namespace bm
