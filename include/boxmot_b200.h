/*
 * boxmot_b200.h -- C ABI of the B200-native association engine (libboxmot_b200.so).
 *
 * The reference (muntherr/boxmot) is pure Python and has no FFI; the entry points
 * below are what a ctypes binding for its per-frame association path binds.  Each
 * declaration cites the reference interface it replaces (paths under the reference
 * tree).  Conventions: every function returns 0 on success and a negative code on
 * failure (message via bm_last_error()); all pointers named d_* are DEVICE pointers
 * owned by the caller; `stream` is a cudaStream_t passed as void*; the library
 * allocates only the opaque engine state at create.  There is no CPU fallback.
 */
#ifndef BOXMOT_B200_H
#define BOXMOT_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* bm_last_error(void);
int bm_version(void);

/* Association functions: boxmot/utils/iou.py:320-346 (AssociationFunction selector). */
enum { BM_ASSO_IOU = 0, BM_ASSO_GIOU = 1, BM_ASSO_DIOU = 2, BM_ASSO_CIOU = 3, BM_ASSO_HMIOU = 4, BM_ASSO_CENTROID = 5 };

/* ---- OC-SORT engine: S independent streams, one reference OcSort instance each.
 * Replaces OcSort.__init__/update (boxmot/trackers/ocsort/ocsort.py:213-244, :246-439),
 * KalmanBoxTracker (:56-192), KalmanFilterXYSR (boxmot/motion/kalman_filters/aabb/
 * xysr_kf.py:137-291) and associate (boxmot/utils/association.py:377-556). */
typedef struct bm_ocsort_config {
    double det_thresh, min_conf, asso_threshold, inertia, q_xy_scaling, q_s_scaling;
    int max_age, min_hits, delta_t, asso_func, use_byte;
    int img_w, img_h;              /* only used by the centroid association function */
} bm_ocsort_config;

/* track_cap / det_cap: per-stream capacities (exceeding them sets an error bit, never
 * truncates silently).  smem_cost_bytes: shared memory reserved per CTA for the cost
 * matrix (larger matrices go to an L2-resident global scratch). */
int bm_ocsort_create(const bm_ocsort_config* cfg, int n_streams, int track_cap, int det_cap,
                     int smem_cost_bytes, void** handle);
int bm_ocsort_destroy(void* handle);
int bm_ocsort_reset(void* handle, void* stream);
/* One frame for all streams.  d_dets [S][det_cap][6] float32 (x1,y1,x2,y2,conf,cls):
 * float32 is the reference's own input precision (basetracker.py:122-128 converts every
 * ndarray).  d_ndets [S] int32.  d_out [S][out_rows][8] float64 rows
 * (x1,y1,x2,y2,id,conf,cls,det_ind) in the reference's row order, out_rows =
 * min(track_cap, det_cap) (a row is only emitted for a track that received a detection
 * this frame), d_nout [S] int32. */
int bm_ocsort_step(void* handle, const float* d_dets, const int* d_ndets, double* d_out,
                   int* d_nout, void* stream);

/* n_frames consecutive frames in one call: d_dets [F][S][det_cap][6], d_ndets [F][S], d_out [F][S][out_rows][8],
 * d_nout [F][S] -- the way the reference's evaluation driver consumes a sequence whose detections are all on disk
 * already (boxmot/engine/val.py:304-354 loops `tracker.update` over the frames of a sequence).  Results are those
 * of n_frames bm_ocsort_step calls.  With stream groups (bm_ocsort_set_groups) every group runs all its frames
 * back to back on its own CUDA stream and the groups are joined once at the end, so a group waiting for the
 * slowest of its streams no longer holds the other groups back frame by frame. */
int bm_ocsort_step_frames(void* handle, const float* d_dets, const int* d_ndets, double* d_out, int* d_nout,
                          int n_frames, void* stream);

/* Step the engine's streams as `groups` independent groups, each on a CUDA stream of its own
 * (forked from and joined back into the caller's stream inside bm_ocsort_step /
 * bm_deepocsort_step): one group's streaming kernels then run under the tail of another group's
 * association kernel.  Results are unchanged (streams are independent, the reference runs one
 * tracker object per video: boxmot/engine/track.py:41-53).  groups <= 1 restores the single
 * chain; per-kernel timing (bm_ocsort_time_kernels) always uses the single chain.  Returns the
 * number of groups in use, < 0 on error. */
int bm_ocsort_set_groups(void* handle, int groups);
/* Synchronises `stream`; returns the OR of all per-stream error bits in *err_bits, the
 * number of live tracks that entered predict since create/reset in *track_updates and
 * the number of lapjv solves in *lap_solves. */
int bm_ocsort_status(void* handle, int* err_bits, unsigned long long* track_updates,
                     unsigned long long* lap_solves, void* stream);
/* Cumulative work counters since create/reset, summed over streams (synchronises
 * `stream`): out4 = {track_updates, lap_solves, matches (Kalman updates with a
 * detection), births}.  Used by bench.py for the algorithmic-bytes figure. */
int bm_ocsort_counters(void* handle, unsigned long long* out4, void* stream);
/* Debug: when d_stamps (device, [S][8] int64) is non-null the association kernel records
 * clock64() at its phase boundaries for every stream; pass NULL to switch it off. */
int bm_ocsort_debug_profile(void* handle, long long* d_stamps);
/* ---- DeepOCSORT engine ------------------------------------------------------------
 * Replaces DeepOcSort.__init__/update (boxmot/trackers/deepocsort/deepocsort.py:265-498),
 * its KalmanBoxTracker (:51-235: update_emb :184-186, apply_affine_correction :191-208 with
 * KalmanFilterXYSR.apply_affine_correction, xysr_kf.py:111-135) and associate with the
 * embedding term (boxmot/utils/association.py:462-474, compute_aw_max_metric :320-374).
 * The engine is the OC-SORT engine with the embedding state added: the handle is accepted by
 * bm_ocsort_destroy / _reset / _status / _counters / _stream_info / _peek / _time_kernels.
 * ReID and CMC are side inputs: the caller passes the detections' embeddings and one 2x3
 * warp per stream.  oc.min_conf and oc.use_byte are ignored (one gate: conf > oc.det_thresh). */
typedef struct bm_deepocsort_config {
    bm_ocsort_config oc;
    double w_association_emb, alpha_fixed_emb, aw_param;
    int embedding_off, aw_off;
    int emb_dim;
} bm_deepocsort_config;
int bm_deepocsort_create(const bm_deepocsort_config* cfg, int n_streams, int track_cap, int det_cap,
                         void** handle);
/* One frame for all streams.  d_dets / d_ndets / d_out / d_nout as bm_ocsort_step.
 * d_embs [S][det_cap][emb_dim] float32, row-aligned with d_dets (may be NULL when
 * embedding_off).  d_warps [S][6] float64, row-major 2x3 camera-motion warp applied to every
 * track before its predict, or NULL (cmc_off / identity).  Any warp is applied: without rotation or
 * shear on the block-form covariance; from a track's first rotating warp on, that track carries a
 * dense 7x7 covariance (apply_affine_correction, deepocsort.py:191-208 + xysr_kf.py:111-135). */
int bm_deepocsort_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                       const double* d_warps, double* d_out, int* d_nout, void* stream);
/* Durations (ms) of the six launches of the last step after bm_ocsort_time_kernels(handle, 1):
 * predict(+CMC), embedding contraction, association, Kalman update, finalize, embedding EMA. */
int bm_deepocsort_kernel_times(void* handle, float* ms6);
/* ---- BoostTrack engine ------------------------------------------------------------
 * Replaces BoostTrack.__init__/update (boxmot/trackers/boosttrack/boosttrack.py:154-336), its
 * KalmanBoxTracker (:45-121) over boosttrack/kalmanfilter.py:29-157, the confidence boosts
 * (dlo_confidence_boost :414-456, duo_confidence_boost :372-412, get_mh_dist_matrix :356-369),
 * filter_outputs (:338-346) and boosttrack/assoc.py (shape_similarity :9-34, MhDist_similarity
 * :37-47, iou_batch :50-66, soft_biou_batch :69-103, match :106-114, linear_assignment :117-153,
 * associate :156-200).  ReID and CMC are side inputs: detection embeddings and one 2x3 warp per
 * stream (any affine warp: camera_update :81-103 rebuilds the state from the warped corners). */
typedef struct bm_boosttrack_config {
    double det_thresh, iou_threshold, min_box_area, aspect_ratio_thresh;
    double lambda_iou, lambda_mhd, lambda_shape, dlo_boost_coef;
    int max_age, min_hits;
    int use_dlo_boost, use_duo_boost, s_sim_corr, use_rich_s, use_sb, use_vt, with_reid;
    int emb_dim;
} bm_boosttrack_config;
int bm_boosttrack_create(const bm_boosttrack_config* cfg, int n_streams, int track_cap, int det_cap,
                         void** handle);
int bm_boosttrack_destroy(void* handle);
int bm_boosttrack_reset(void* handle, void* stream);
/* One frame for all streams; arguments as bm_deepocsort_step (d_embs may be NULL without
 * with_reid, d_warps NULL = no camera_update this frame).  Output rows are in the reference's
 * order (tracker list order); the reported confidence is the BOOSTED detection confidence. */
int bm_boosttrack_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                       const double* d_warps, double* d_out, int* d_nout, void* stream);
/* per_class mode of BoostTrack: the decorator calls update once per class id, but BoostTrack keeps its tracks in
 * self.trackers (boosttrack.py:185), which the decorator does not swap -- every class call predicts, matches and ages
 * the ONE pool with that class's detections, at the same frame_count.  new_frame = 1 records the frame counter before
 * the frame's first class call, new_frame = 0 rewinds it before each of the others. */
int bm_boosttrack_hold_frame(void* handle, int new_frame, void* stream);
/* Stream groups for the BoostTrack engine; see bm_ocsort_set_groups. */
int bm_boosttrack_set_groups(void* handle, int groups);
/* Synchronises `stream`: OR of the per-stream error bits and cumulative {track_updates,
 * lap_solves, matches, births}. */
int bm_boosttrack_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream);
/* Per-launch CUDA-event timing of the following steps: predict, boosts, embedding contraction,
 * association, finalize, embedding EMA (ms). */
int bm_boosttrack_time_kernels(void* handle, int enable);
int bm_boosttrack_kernel_times(void* handle, float* ms6);
/* Test hook: Kalman means [n][8], covariances [n][8][8] and ids of one stream in list order. */
int bm_boosttrack_peek(void* handle, int stream_index, int* n_tracks, double* h_x, double* h_P,
                       int* h_ids, void* stream);
/* ---- ByteTrack engine -------------------------------------------------------------
 * Replaces ByteTrack.__init__/update (boxmot/trackers/bytetrack/bytetrack.py:132-156,
 * :158-302), STrack (:14-116), joint/sub/remove_duplicate_stracks (:308-346),
 * KalmanFilterXYAH (boxmot/motion/kalman_filters/aabb/xyah_kf.py, base_kalman_filter.py:
 * 43-155), iou_distance / fuse_score / linear_assignment (boxmot/utils/matching.py:184-227,
 * :488-560, :30-108).  Same buffer conventions as bm_ocsort_step: d_dets [S][det_cap][6]
 * float32, d_ndets [S], d_out [S][min(track_cap,det_cap)][8] float64 rows
 * (x1,y1,x2,y2,id,conf,cls,det_ind) in active_tracks order, d_nout [S]. */
typedef struct bm_bytetrack_config {
    double min_conf, track_thresh, match_thresh;
    int track_buffer, frame_rate;
} bm_bytetrack_config;
int bm_bytetrack_create(const bm_bytetrack_config* cfg, int n_streams, int track_cap, int det_cap,
                        void** handle);
int bm_bytetrack_destroy(void* handle);
int bm_bytetrack_reset(void* handle, void* stream);
int bm_bytetrack_step(void* handle, const float* d_dets, const int* d_ndets, double* d_out,
                      int* d_nout, void* stream);
/* per_class mode of the BYTE family (also for BoT-SORT handles): the reference's per_class_decorator
 * (boxmot/trackers/basetracker.py:154-201) calls update once per class id 0 .. nr_classes-1 per frame, each time with
 * per_class_active_tracks[cls] as the tracker's active_tracks and frame_count reset to the frame's starting value;
 * lost_stracks / removed_stracks stay ONE list for all classes (bytetrack.py:143-144, botsort.py:68-69), which
 * couples the classes.  This call makes class `cls`'s active list the engine's (saving the current one) and, unless
 * new_frame, rewinds the frame counter; the caller then steps the class's detections (det_ind relative to them).
 * new_frame = 1 on the frame's first class.  n_classes is fixed by the first call.  Asynchronous on `stream`. */
int bm_bytetrack_select_class(void* handle, int n_classes, int cls, int new_frame, void* stream);
/* out4 = {pool tracks predicted, assignments solved, matches, births}, summed over streams;
 * err_bits = OR of the per-stream error bits.  Synchronises the stream. */
int bm_bytetrack_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream);
/* Host copy of one stream: active tracks then lost tracks in list order; h_mean [n][8],
 * h_cov [n][64] (dense), h_ids [n]; any pointer may be NULL. */
int bm_bytetrack_peek(void* handle, int stream_index, int* n_active, int* n_lost, double* h_mean,
                      double* h_cov, int* h_ids, void* stream);

/* ---- BoT-SORT engine ---------------------------------------------------------------
 * Replaces BotSort.__init__/update and its stages (boxmot/trackers/botsort/botsort.py:49-92,
 * :94-411), STrack incl. update_features / update_cls / multi_predict / multi_gmc
 * (botsort_track.py:10-159), botsort_utils.py:10-81, KalmanFilterXYWH, and iou_distance /
 * fuse_score / embedding_distance / linear_assignment (boxmot/utils/matching.py:184-227,
 * :488-560, :230-321, :30-108).  ReID and CMC are side inputs: d_embs [S][det_cap][emb_dim]
 * float32 rows aligned with d_dets (NULL when with_reid = 0), d_warps [S][6] float64 row-major
 * 2x3 camera-motion warps or NULL for the identity (any warp: a track switches to a dense 8x8
 * covariance at its first warp with rotation / shear, multi_gmc, botsort_track.py:91-104).  Other
 * buffers as bm_bytetrack_step.  One frame = three
 * launches: embedding kernel (feature preparation + cosine-distance contraction), the step,
 * the feature EMA. */
typedef struct bm_botsort_config {
    double track_high_thresh, track_low_thresh, new_track_thresh, match_thresh, proximity_thresh,
        appearance_thresh;
    int track_buffer, frame_rate, fuse_first_associate, with_reid, emb_dim;
} bm_botsort_config;
int bm_botsort_create(const bm_botsort_config* cfg, int n_streams, int track_cap, int det_cap,
                      void** handle);
int bm_botsort_destroy(void* handle);
int bm_botsort_reset(void* handle, void* stream);
int bm_botsort_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                    const double* d_warps, double* d_out, int* d_nout, void* stream);
/* Stream groups for the BoT-SORT engine; see bm_ocsort_set_groups. */
int bm_botsort_set_groups(void* handle, int groups);
int bm_botsort_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream);
int bm_botsort_peek(void* handle, int stream_index, int* n_active, int* n_lost, double* h_mean,
                    double* h_cov, int* h_ids, void* stream);

/* ---- StrongSORT engine ---------------------------------------------------------------
 * Replaces StrongSort.__init__/update/_compute_detection_quality/_format_outputs
 * (boxmot/trackers/strongsort/strongsort.py:43-113, :115-205, :301-327, :329-359), Tracker
 * (sort/tracker.py:102-111 predict, :118-181 update, :183-281 _enhanced_match, :283-298
 * _apply_id_preservation_weighting, :300-351 _attempt_id_recovery, :395-410 _initiate_track),
 * matching_cascade / min_cost_matching / gate_cost_matrix and helpers / NearestNeighborDistanceMetric
 * (sort/linear_assignment.py:14-87, :90-171, :174-379, :499-618), iou_cost (sort/iou_matching.py:49-87),
 * Track (sort/track.py:78-317) and scipy.optimize.linear_sum_assignment as called at
 * linear_assignment.py:70 (emulated step for step, so ties resolve as scipy's do).
 * The fork runs only in one configuration: R4 (unmatched_tracks_3 initialised), the
 * occlusion handler and crowd detection off (R5), and -- for anything to be confirmed -- tracks born
 * Confirmed as under GITHUB_ACTIONS=true (M1: born_confirmed = 1).  ReID and CMC are side inputs:
 * d_embs [S][det_cap][emb_dim] float32 (required), d_warps [S][6] float64 row-major 2x3 warps or
 * NULL for the identity (camera_update runs either way, strongsort.py:168-172; any affine warp).
 * Output rows are the first 8 of the fork's 10 columns, in tracker.tracks order.
 * pool_cap: feature vectors stored per stream (each distinct vector once: the <= 10 features of
 * every track and whatever its sample list of nn_budget entries still references). */
typedef struct bm_strongsort_config {
    double min_conf, max_cos_dist, max_iou_dist, mc_lambda, ema_alpha, conf_thresh_high, conf_thresh_low,
        id_preservation_weight;
    int max_age, n_init, nn_budget, born_confirmed, emb_dim;
} bm_strongsort_config;
int bm_strongsort_create(const bm_strongsort_config* cfg, int n_streams, int track_cap, int det_cap, int pool_cap,
                         void** handle);
int bm_strongsort_destroy(void* handle);
/* scipy.optimize.linear_sum_assignment (StrongSORT's solver, sort/linear_assignment.py:70) for a batch
 * of stored matrices, emulated step for step (Crouse's shortest augmenting path as in scipy's
 * rectangular_lsap.cpp: rows in order, the `remaining` column list with its swap-removal, ties to an
 * unassigned column, transposition when nc < nr), so that the result equals scipy's also when the optimum
 * is not unique.  d_cost [B][nr_max][nc_max] float64, d_x [B][nr_max] = column of each row or -1.
 * d_work: caller-owned scratch of at least B * 48 * max(nr_max, nc_max) bytes. */
int bm_linear_sum_assignment_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                                   const int* d_nc, int* d_x, void* d_work, void* stream);
/* The same assignment for the shape of StrongSORT's IoU stage (sort/tracker.py:257-275): many rows, few columns,
 * and most rows "far" -- d_far_rows [B][nr_max] != 0 promises that the row equals `cstar` in every column (tracks
 * with time_since_update > 1 after the clamp of min_cost_matching).  The far rows are then handled implicitly
 * (never touched by the augmenting-path sweeps) with the SAME result as the full emulation, scipy's tie-breaking
 * included; problems outside the fast path's limits are solved on the full matrix.  d_used_far [B] (may be NULL)
 * reports which path ran.  d_work as bm_linear_sum_assignment_batch. */
int bm_linear_sum_assignment_far_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                                       const int* d_nc, const int* d_far_rows, double cstar, int* d_x, int* d_used_far,
                                       void* d_work, void* stream);
int bm_strongsort_reset(void* handle, void* stream);
/* One frame for all streams; buffers as bm_botsort_step, out_rows = det_cap.  Error bits (see
 * bm_strongsort_counters): 1 track_cap, 2 det_cap, 8 pool_cap exceeded, 16 assignment guard. */
int bm_strongsort_step(void* handle, const float* d_dets, const int* d_ndets, const float* d_embs,
                       const double* d_warps, double* d_out, int* d_nout, void* stream);
/* out4 = {tracks entering predict, assignments solved, matches (Track.update calls), births}. */
int bm_strongsort_counters(void* handle, int* err_bits, unsigned long long* out4, void* stream);
/* Capacity planning: over all streams, the largest number of track slots in use (tracks + lost buffer),
 * of pool slots in use and of vectors referenced by a sample list (rows of the contraction) right now. */
int bm_strongsort_usage(void* handle, int* max_tracks, int* max_pool_used, int* max_live, void* stream);
/* Stream groups for the StrongSORT engine; see bm_ocsort_set_groups.  Results do not depend on it; measured at
 * configs[3] it does not pay (one chain 6.58 ms, 2-8 groups 6.9 ms), so the default is one chain. */
int bm_strongsort_set_groups(void* handle, int groups);
/* Debug: host copy of one stream's scratch (which = 0: the cascade's integer work area, 1: the cost matrix
 * [track_cap][det_cap] float64 by list position and sorted detection, 2: the sorted detection table). */
int bm_strongsort_debug_copy(void* handle, int which, int stream_index, void* h_dst, size_t nbytes, void* stream);
/* Per-launch CUDA-event timing: prepare+predict, bank contraction, cost matrix, cascade, updates, finalize (the
 * finalize, sample-list and live-list launches together) (ms). */
int bm_strongsort_time_kernels(void* handle, int enable);
int bm_strongsort_kernel_times(void* handle, float* ms6);
/* Test hook: tracks of one stream in list order: means [n][8], covariances [n][64], table rows
 * [n][42] (bm_strongsort_track's columns), ids, sample-list lengths; any pointer may be NULL. */
int bm_strongsort_peek(void* handle, int stream_index, int* n_tracks, int* n_lost, double* h_mean, double* h_cov,
                       double* h_tab, int* h_ids, int* h_nbank, void* stream);

/* Per-kernel timing of the frame step (measurement aid for bench.py's roofline): when
 * enabled, bm_ocsort_step brackets its four launches (predict, assoc, update, finalize) with
 * CUDA events on the caller's stream; bm_ocsort_kernel_times waits for the last timed step and
 * returns the four durations in milliseconds. */
int bm_ocsort_time_kernels(void* handle, int enable);
int bm_ocsort_kernel_times(void* handle, float* ms4);

/* Per-stream live-track count and number of tracks created so far (h_* [S], either may be
 * NULL).  KalmanBoxTracker.count (ocsort.py:60,79-80) is ONE counter per tracker object in the
 * reference: the per_class adapter (basetracker.py:154-201) renumbers the per-class ids with
 * the creation counts of every class.  Synchronises the stream. */
int bm_ocsort_stream_info(void* handle, int* h_n_tracks, int* h_next_id, void* stream);

/* Debug / test access: copy one stream's list length and per-track Kalman state
 * (list order) to host buffers x[track_cap*7], P[track_cap*49], ids[track_cap]. */
int bm_ocsort_peek(void* handle, int stream_index, int* n_tracks, double* h_x, double* h_P,
                   int* h_ids, void* stream);

/* The step functions' padded rows d_out [S][out_rows][8] / d_nout [S] as ONE dense ragged buffer:
 * stream s owns rows [offsets[s], offsets[s + 1]) of rows_dst (offsets_dst has S + 1 entries, the last
 * one the total).  What BaseTracker.update returns per stream (basetracker.py:67-81: an (M, 8) array) for
 * all streams at once.  rows_dst / offsets_dst may be device memory or MAPPED PINNED HOST memory: then
 * the kernel's stores are the device-to-host transfer, 64 bytes per emitted row instead of the padded
 * array.  rows_dst must hold sum(nout) rows (S * out_rows always suffices). */
int bm_pack_rows(const double* d_out, const int* d_nout, int n_streams, int out_rows, double* rows_dst,
                 int* offsets_dst, void* stream);

/* convert_to_mot_format (boxmot/engine/utils.py:101-146, the ndarray branch) for all streams of a frame,
 * fused with bm_pack_rows' ragged packing: rows_dst [sum(nout)][9] float64 = frame_idx, id, left, top,
 * width, height (xyxy2ltwh, rounded half-to-even as np.round, cast to int32), 1, class, confidence --
 * what write_mot_results (:149-172) prints with "%d,%d,%d,%d,%d,%d,%d,%d,%.6f".  Destinations as
 * bm_pack_rows (device or mapped pinned host memory). */
int bm_pack_mot_rows(const double* d_out, const int* d_nout, int n_streams, int out_rows, int frame_idx,
                     double* rows_dst, int* offsets_dst, void* stream);

/* ---- unit-testable kernels ---------------------------------------------------- */

/* lap.lapjv(cost, extend_cost=True, cost_limit) for a batch of problems (call sites:
 * boxmot/utils/matching.py:54-55, association.py:50-51,109, boosttrack/assoc.py:113).
 * d_cost [B][nr_max][nc_max] float64, per-problem sizes d_nr/d_nc, cost_limit = +inf
 * for "no limit".  d_x [B][nr_max], d_y [B][nc_max] (-1 = unassigned). */
int bm_lap_solve_batch(const double* d_cost, int n_problems, int nr_max, int nc_max,
                       const int* d_nr, const int* d_nc, double cost_limit, int* d_x, int* d_y,
                       void* stream);

/* Same contract and arguments, solved by the engine's fast exact solver (real rows only,
 * implicit null option; csrc/lap_rect.cuh).  Equal to bm_lap_solve_batch whenever the
 * optimum is unique; under exact cost ties it may return a different optimal solution. */
int bm_lap_rect_batch(const double* d_cost, int n_problems, int nr_max, int nc_max,
                      const int* d_nr, const int* d_nc, double cost_limit, int* d_x, int* d_y,
                      void* stream);

/* AssociationFunction.*_batch (boxmot/utils/iou.py:50-307): d_a [A][4], d_b [B][4]
 * xyxy float64 -> d_out [A][B] float64. */
int bm_iou_matrix(int asso_func, const double* d_a, int na, const double* d_b, int nb,
                  double img_w, double img_h, double* d_out, void* stream);

/* KalmanFilterXYSR.predict / update with OC-SORT constants (xysr_kf.py:137-175,
 * :256-291; ocsort.py:83-113): d_x [N][7], d_P [N][49] updated in place, d_z [N][4]. */
int bm_xysr_predict(double* d_x, double* d_P, int n, double q_xy, double q_s, void* stream);
int bm_xysr_update(double* d_x, double* d_P, const double* d_z, int n, void* stream);

/* The same filter in the engine's block form (csrc/kf_xysr.cuh): d_Pb [N][13] holds, for
 * b = 0,1,2, P[b][b], P[b][b+4], P[b+4][b], P[b+4][b+4] and then P[3][3] -- the only
 * entries that are ever non-zero with the OC-SORT constants.  Bit-identical to
 * bm_xysr_predict / bm_xysr_update on such covariances. */
int bm_xysrb_predict(double* d_x, double* d_Pb, int n, double q_xy, double q_s, void* stream);
int bm_xysrb_update(double* d_x, double* d_Pb, const double* d_z, int n, void* stream);

/* Constant-velocity box Kalman filters, batched (unit-testable form of what the ByteTrack
 * engine runs per track).  kind 0 = KalmanFilterXYAH (boxmot/motion/kalman_filters/aabb/
 * xyah_kf.py), 1 = KalmanFilterXYWH (xywh_kf.py); op 0 = initiate (base_kalman_filter.py:43-53;
 * reads d_z, writes mean/cov), 1 = predict (:62-77), 2 = update (:129-155 with project :86-109;
 * d_conf [n] is the NSA confidence, NULL = 0), 3 = multi_predict (:111-127: the same update, but
 * np.dot(np.dot(F, P), F^T) where predict's multi_dot evaluates F (P F^T)).  Bit-identical to the
 * reference's numpy / scipy arithmetic (both off-diagonal entries of each block are carried).
 * d_mean [n][8], d_cov [n][8][8] dense, d_z [n][4].  Covariances must have the filter's own
 * sparsity (only (pos_i, vel_i) coupled), which initiate/predict/update preserve. */
int bm_kf_cv8(int kind, int op, double* d_mean, double* d_cov, const double* d_z, const double* d_conf,
              int n, void* stream);
/* gating_distance (base_kalman_filter.py:166-194) of every (track, measurement) pair:
 * d_out [n_tracks][n_meas]; gaussian = 0 -> squared Mahalanobis ("maha"), 1 -> "gaussian". */
int bm_kf_cv8_gating(int kind, const double* d_mean, const double* d_cov, int n_tracks, const double* d_meas,
                     int n_meas, int only_position, int gaussian, double* d_out, void* stream);

/* StrongSORT's appearance metric NearestNeighborDistanceMetric.distance with the cosine metric
 * (boxmot/trackers/strongsort/sort/linear_assignment.py:595-618, _nn_cosine_distance :468-497,
 * _cosine_distance :382-413): d_out [n_tracks][n_feats] float32 = min over the track's stored
 * samples of 1 - clip(<f / (|f| + 1e-8), g / (|g| + 1e-8)>, -1, 1), or 1e5 (INFTY_COST, :11) for
 * a track without samples.  d_bank [n_tracks][bank_cap][emb_dim] float32 with d_nb [n_tracks]
 * live samples each; d_feats [n_feats][emb_dim].  One tcgen05 contraction per track (3xTF32)
 * with the minimum taken in the epilogue.  d_work: caller-owned scratch of at least
 * 4 * (max(bank_cap, n_feats) + n_tracks * bank_cap + n_feats) bytes. */
int bm_nn_cosine_distance(const float* d_bank, const int* d_nb, int n_tracks, int bank_cap, const float* d_feats,
                          int n_feats, int emb_dim, float* d_out, void* d_work, void* stream);

/* StrongSORT's min_cost_matching on given cost matrices (strongsort/sort/linear_assignment.py:
 * 14-87): entries > max_distance are set to max_distance + 1e-5 (:67), scipy's rectangular
 * linear_sum_assignment (:70) is replaced by the exact rectangular solver, and assigned pairs
 * whose cost is > max_distance are rejected (:82).  Arguments as bm_lap_rect_batch; d_x / d_y
 * hold the accepted pairs (-1 = unmatched). */
int bm_min_cost_matching_batch(const double* d_cost, int n_problems, int nr_max, int nc_max, const int* d_nr,
                               const int* d_nc, double max_distance, int* d_x, int* d_y, void* stream);

/* StrongSORT cost post-processing, one entry per (track, detection) pair.  d_trk [n_tracks][13]
 * doubles per track: age, hits, time_since_update, motion_consistency, appearance_consistency,
 * quality_score, high_conf_streak, low_conf_streak, has_velocity (velocity_history non-empty),
 * velocity_history[-1] x, y, conf, cls (strongsort/sort/track.py:86-142).
 *
 * bm_strongsort_gated_metric: what Tracker._enhanced_match.gated_metric (sort/tracker.py:186-206)
 * does to the matrix metric.distance returned (bm_nn_cosine_distance): gate_cost_matrix
 * (sort/linear_assignment.py:174-248: squared Mahalanobis gate at chi2inv95[4] -> 1e5, motion cost
 * :288-307, adaptive lambda :310-329, blend :241-243, track adjustments :332-352) and the
 * id-preservation weighting (sort/tracker.py:283-298).  d_cost [n_tracks][n_dets] fp64 in place;
 * d_mean [n_tracks][8], d_cov [n_tracks][8][8] (KalmanFilterXYAH state), d_meas [n_dets][4] xyah.
 *
 * bm_strongsort_enhance_cost: _enhance_cost_matrix (sort/linear_assignment.py:251-273) applied by
 * min_cost_matching before the clamp; d_det [n_dets][3] = quality_score, conf, cls.
 *
 * bm_strongsort_iou_cost: iou_cost (sort/iou_matching.py:49-87) of Track.to_tlwh() against
 * d_tlwh [n_dets][4]; rows of tracks with time_since_update > 1 are 1e5. */
int bm_strongsort_gated_metric(double* d_cost, const double* d_mean, const double* d_cov, const double* d_trk,
                               int n_tracks, const double* d_meas, int n_dets, double mc_lambda, double id_weight,
                               void* stream);
int bm_strongsort_enhance_cost(double* d_cost, const double* d_trk, int n_tracks, const double* d_det, int n_dets,
                               void* stream);
int bm_strongsort_iou_cost(const double* d_mean, const double* d_trk, int n_tracks, const double* d_tlwh, int n_dets,
                           double* d_out, void* stream);

/* Feature EMA of StrongSORT's Track.update (strongsort/sort/track.py:238-268 with
 * _calculate_adaptive_ema_alpha :319-335) for n updated tracks: the detection feature is normalised
 * (|f| + 1e-8), compared with the track's newest stored feature d_last [n][emb_dim] (cosine
 * similarity, written to d_similarity [n] when not NULL), folded into appearance_consistency
 * (0.9 / 0.1, d_appearance_consistency [n] in place), and blended with the adaptive alpha =
 * clip(ema_alpha * conf_factor * appearance_factor, 0.1, 0.95), renormalised: d_out [n][emb_dim]
 * is the feature Track.update appends.  d_has_last [n] = 0 for a track without features (first
 * feature: the normalised detection feature, consistency untouched).  fp64 arithmetic, fp32 storage. */
int bm_strongsort_feature_update(const float* d_last, const int* d_has_last, const float* d_feat, const double* d_conf,
                                 int n, int emb_dim, double ema_alpha, double* d_appearance_consistency,
                                 double* d_similarity, float* d_out, void* stream);

/* StrongSORT's per-track state machine (strongsort/sort/track.py) on a flat table, one row of 42
 * doubles per track: columns 0..12 as d_trk of bm_strongsort_gated_metric (column 8 =
 * len(velocity_history) capped at 10, 9..10 = velocity_history[-1]), 13 state (1 tentative, 2
 * confirmed, 3 deleted), 14 stability_score, 15 confirmed_detections, 16 missed_detections,
 * 17 len(position_history) capped at 10, 18..19 position_history[-1], 20 len(confidence_history),
 * 21..40 confidence_history (oldest first, at most 20), 41 det_ind.  d_mean [n][8], d_cov [n][8][8].
 *   op 0  Track.predict (:195-219): Kalman predict, age / time_since_update, velocity and position
 *         history, motion consistency (:379-399);
 *   op 1  Track.update (:221-301) without its feature step (run bm_strongsort_feature_update first;
 *         its appearance consistency goes to column 4): NSA Kalman update with the detection
 *         confidence, confidence history, streaks, hits, quality score (:337-359), stability score
 *         (:361-377), tentative -> confirmed (:298-301); d_det [n][7] = x, y, a, h, conf, cls, det_ind;
 *   op 2  Track.mark_missed (:303-317): quality-dependent deletion threshold;
 *   op 3  Track.camera_update (:173-189): the corners of to_tlbr() through the 2x3 warp in d_det [n][7]
 *         (first six entries, row-major), mean[:4] rebuilt, motion consistency; no covariance change. */
int bm_strongsort_track(int op, double* d_mean, double* d_cov, double* d_trk, const double* d_det, int n, int n_init,
                        int max_age, void* stream);

/* embedding_distance (boxmot/utils/matching.py:230-321 with its defaults): d_out [na][nb] =
 * max(0, 1 - <a_i, b_j> / (|a_i| |b_j|)) for float32 feature rows d_a [na][emb_dim],
 * d_b [nb][emb_dim].  mode 0 = tcgen05 tensor-core contraction (3xTF32, fp32 accumulation in
 * TMEM), 1 = fp32 FFMA.  d_work: caller-owned scratch of at least 8 * (na + nb) bytes. */
int bm_cosine_distance(const float* d_a, int na, const float* d_b, int nb, int emb_dim, float* d_out,
                       void* d_work, int mode, void* stream);

/* bm_cosine_distance for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE
 * REFERENCE.  Same result as bm_cosine_distance
 * (same kernels), but the row-index / norm scratch lives in a file-scope static __device__ array
 * of the library instead of a caller buffer: 8 * BM_COSINE_NOALLOC_MAX_ROWS = 128 KiB, held by
 * every GPU that loads the library until the process exits.  The call performs no cudaMalloc /
 * cudaFree and calls no allocating library routine; the first call on each device resolves the
 * array's device address (cudaGetSymbolAddress) and sets the kernel's shared-memory attribute, so
 * make that first call outside stream capture.  na + nb > BM_COSINE_NOALLOC_MAX_ROWS is refused
 * with -4 and a message in bm_last_error().  All calls on a device SHARE the one array: it keeps
 * the previous call's contents (every entry a call reads is rewritten by that call first, so
 * nothing needs clearing), and two calls that overlap on the GPU (different CUDA streams, or
 * graph branches) race on it -- serialise them on one stream.  Inputs and output are still the
 * caller's: d_a [na][emb_dim], d_b [nb][emb_dim] float32, d_out [na][nb] float32, mode as
 * bm_cosine_distance. */
#define BM_COSINE_NOALLOC_MAX_ROWS 16384
int bm_cosine_distance_noalloc(const float* d_a, int na, const float* d_b, int nb, int emb_dim, float* d_out,
                               int mode, void* stream);

#ifdef __cplusplus
}
#endif
#endif
