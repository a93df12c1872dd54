#!/usr/bin/env python
"""Headline benchmark: OC-SORT association hot path on 1024 concurrent synthetic streams
(BASELINE.json configs[1]: 100 objects/frame, 300 frames, IoU + OCM velocity cost, lapjv).

    python bench.py --gpus N --steps K --warmup W            # B200 engine (default arm)
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port of
                                                             # the reference path on host cores

A "step" is ONE frame advanced for every stream of the batch (one launch of the fused
frame-step kernel).  `value` = track-updates/s (live tracks entering predict, summed over
streams and frames) with inputs resident in HBM; `e2e` = the same metric
through the host-buffer API with H2D/D2H inside the timed region.  Prints ONE JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

OCSORT_YAML = dict(min_conf=0.1, det_thresh=0.6, max_age=30, min_hits=3, asso_threshold=0.3,
                   delta_t=3, asso_func="iou", inertia=0.1, use_byte=False, Q_xy_scaling=0.01,
                   Q_s_scaling=0.0001)     # boxmot/configs/trackers/ocsort.yaml defaults
BOTSORT_YAML = dict(track_high_thresh=0.6, track_low_thresh=0.1, new_track_thresh=0.7, track_buffer=30,
                    match_thresh=0.8, proximity_thresh=0.5, appearance_thresh=0.25, frame_rate=30,
                    fuse_first_associate=False, with_reid=True)   # boxmot/configs/trackers/botsort.yaml
BYTETRACK_YAML = dict(min_conf=0.1, track_thresh=0.6, match_thresh=0.9, track_buffer=30,
                      frame_rate=30)   # boxmot/configs/trackers/bytetrack.yaml defaults
DEEPOCSORT_YAML = dict(det_thresh=0.5, max_age=30, min_hits=3, iou_threshold=0.3, delta_t=3, asso_func="iou",
                       inertia=0.2, w_association_emb=0.75, alpha_fixed_emb=0.95, aw_param=0.5, embedding_off=False,
                       cmc_off=False, aw_off=False, Q_xy_scaling=0.01,
                       Q_s_scaling=0.0001)   # boxmot/configs/trackers/deepocsort.yaml defaults
BOOSTTRACK_YAML = dict(max_age=60, min_hits=3, det_thresh=0.6, iou_threshold=0.3, use_ecc=True, min_box_area=10,
                       aspect_ratio_thresh=1.6, lambda_iou=0.5, lambda_mhd=0.25, lambda_shape=0.25,
                       use_dlo_boost=True, use_duo_boost=True, dlo_boost_coef=0.65, s_sim_corr=False, use_rich_s=True,
                       use_sb=True, use_vt=True, with_reid=True)   # boxmot/configs/trackers/boosttrack.yaml defaults
STRONGSORT_ARGS = dict(min_conf=0.1, max_cos_dist=0.15, max_iou_dist=0.7, max_age=50, n_init=2, nn_budget=50,
                       mc_lambda=0.995, ema_alpha=0.9, conf_thresh_high=0.7, conf_thresh_low=0.3,
                       id_preservation_weight=0.1)   # StrongSort.__init__ defaults, 50-deep sample bank (BASELINE configs[3])
EMB_TRACKERS = ("botsort", "deepocsort", "boosttrack", "strongsort")
KERNEL_NAMES = {
    "ocsort": ("ocs_predict_kernel", "ocs_assoc_kernel", "ocs_update_kernel", "ocs_finalize_kernel"),
    "deepocsort": ("ocs_predict_kernel<deep>", "cosine_tc_kernel<DeepDot>", "ocs_assoc_kernel<deep>",
                   "ocs_update_kernel", "ocs_finalize_kernel", "docs_emb_kernel"),
    "boosttrack": ("bt_predict_kernel", "bt_boost_kernel", "cosine_tc_kernel<BoostDot>",
                   "bt_assoc_kernel+bt_update_kernel", "bt_finalize_kernel", "bt_emb_kernel"),
    "strongsort": ("ss_prep_kernel", "cosine_tc_kernel<SsBank>", "ss_cost_kernel", "ss_match_kernel",
                   "ss_update_kernel", "ss_finalize_kernel"),
}
METRIC = "track-updates/sec"
L2_BYTES = 126 * 1024 * 1024


_REAL_STDOUT = None


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="default 290 (botsort: 40)")
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tracker", default="ocsort", choices=["ocsort", "bytetrack", "botsort", "deepocsort", "boosttrack", "strongsort"],
                    help="ocsort = the headline workload (BASELINE configs[1]); bytetrack = the same "
                         "synthetic streams through the ByteTrack engine (the tracker of configs[0]); "
                         "botsort = BASELINE configs[2] (200 dets/frame, 512-d embeddings); deepocsort / boosttrack = the "
                         "trackers of configs[4] on the configs[1] streams with 512-d embeddings (their engines keep "
                         "per-column operands in shared memory: crowd shapes need --track-cap <= 768)")
    ap.add_argument("--emb-dim", type=int, default=None, help="embedding trackers: default 512 (strongsort: 768)")
    ap.add_argument("--pool-cap", type=int, default=None, help="strongsort: feature vectors stored per stream")
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--objects", type=int, default=None, help="default 100 (botsort: 200)")
    ap.add_argument("--det-cap", type=int, default=None, help="default 128 (botsort: 256)")
    ap.add_argument("--track-cap", type=int, default=None, help="default 256 (botsort: 512)")
    ap.add_argument("--groups", type=int, default=None,
                    help="independent stream groups stepped on CUDA streams of their own (engine.set_stream_groups)")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between steps")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sub-configs", dest="sub_configs", action="store_false",
                    help="headline run only: skip the short sub-runs of BASELINE configs[0], [2], [3], [4]")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-streams", type=int, default=4, help="cpu_baseline sample size")
    ap.add_argument("--ref-streams-per-core", type=int, default=1)
    a = ap.parse_args()
    bot = a.tracker == "botsort"
    emb = a.tracker in EMB_TRACKERS
    if a.tracker == "strongsort":
        # BASELINE configs[3]: 768-d embeddings, 50-deep sample bank.  Born-Confirmed tracks (the fork's only
        # runnable mode) make every low-confidence detection a track that lives max_age frames: ~12 live
        # tracks per object at steady state, hence the track / pool capacities and the long warm-up.
        for name, v in (("steps", 30), ("warmup", 90), ("objects", 100), ("det_cap", 128), ("track_cap", 2048),
                        ("pool_cap", 8192), ("emb_dim", 768), ("streams", 512)):       # 4096 streams / 8 GPUs
            if getattr(a, name) is None or (name in ("warmup", "streams") and getattr(a, name) == ap.get_default(name)):
                setattr(a, name, v)
    if a.emb_dim is None:
        a.emb_dim = 512
    for name, dflt, dbot in (("steps", 290, 40), ("objects", 100, 200), ("det_cap", 128, 256), ("track_cap", 256, 384)):     # BoT-SORT: pool ~ 160 tracks at 200 objects/frame; 384 slots = 4 step CTAs per SM (512: 3)
        if getattr(a, name) is None:
            setattr(a, name, dbot if bot else (40 if (emb and name == "steps") else dflt))
    if not emb:
        a.emb_dim = 0
    return a


DEFAULT_GROUPS = {"strongsort": 1, "ocsort": 2, "deepocsort": 4, "boosttrack": 4, "botsort": 8}      # tracker -> stream groups per engine step (measured optimum; 1 = one chain of launches)


# --------------------------------------------------------------------------- inputs
_SHARED_EMBS = None      # [F, S, dcap, E] float32 in an anonymous shared mapping (fork workers fill it)


def _gen_one(a):
    from boxmot_b200.synth import make_stream
    sid, k, f, dcap = a[:4]
    if len(a) > 4 and a[4]:
        fr, em = make_stream(sid, k, f, emb_dim=a[4])
        for i, e in enumerate(em):
            _SHARED_EMBS[i, a[5], : len(e)] = e
    else:
        fr = make_stream(sid, k, f)
    d = np.zeros((f, dcap, 6), np.float32)
    n = np.zeros((f,), np.int32)
    for i, x in enumerate(fr):
        if len(x) > dcap:
            raise ValueError("frame with %d detections > det_cap %d" % (len(x), dcap))
        d[i, : len(x)] = x
        n[i] = len(x)
    return d, n


def gen_inputs(first_stream, n_streams, k, n_frames, dcap, workers, emb_dim=0):
    """dets [F,S,dcap,6] float32, ndets [F,S] int32 (+ embs [F,S,dcap,E] float32 when emb_dim);
    generated before CUDA is touched."""
    import mmap
    import multiprocessing as mp
    global _SHARED_EMBS
    dets = np.zeros((n_frames, n_streams, dcap, 6), np.float32)
    nd = np.zeros((n_frames, n_streams), np.int32)
    if emb_dim:
        shape = (n_frames, n_streams, dcap, emb_dim)
        buf = mmap.mmap(-1, int(np.prod(shape)) * 4)
        _SHARED_EMBS = np.frombuffer(buf, dtype=np.float32).reshape(shape)
    jobs = [(first_stream + s, k, n_frames, dcap, emb_dim, s) for s in range(n_streams)]
    if workers > 1:
        with mp.get_context("fork").Pool(workers) as pool:
            for s, (d, n) in enumerate(pool.imap(_gen_one, jobs, chunksize=8)):
                dets[:, s], nd[:, s] = d, n
    else:
        for s, j in enumerate(jobs):
            dets[:, s], nd[:, s] = _gen_one(j)
    if emb_dim:
        return dets, nd, _SHARED_EMBS
    return dets, nd


# --------------------------------------------------------------------------- CPU arms
def _cpu_worker(a):
    """Runs whole streams through the oracle port; returns (timed seconds, timed updates)."""
    os.environ["OMP_NUM_THREADS"] = "1"
    from boxmot_b200.synth import make_stream
    sids, k, warm, steps, tracker = a[:5]
    emb_dim = a[5] if len(a) > 5 else 0
    if tracker == "botsort":
        from oracle.botsort import BotSortOracle
        make = lambda: BotSortOracle(**BOTSORT_YAML)
    elif tracker == "deepocsort":
        from oracle.deepocsort import DeepOcSortOracle
        make = lambda: DeepOcSortOracle(**DEEPOCSORT_YAML)
    elif tracker == "boosttrack":
        from oracle.boosttrack import BoostTrackOracle
        make = lambda: BoostTrackOracle(**BOOSTTRACK_YAML)
    elif tracker == "strongsort":
        from oracle.strongsort_tracker import StrongSortOracle
        make = lambda: StrongSortOracle(**STRONGSORT_ARGS)
    elif tracker == "bytetrack":
        from oracle.bytetrack import ByteTrackOracle
        make = lambda: ByteTrackOracle(**BYTETRACK_YAML)
    else:
        from oracle.ocsort import OcSortOracle
        make = lambda: OcSortOracle(**OCSORT_YAML)
    tot_t, tot_u = 0.0, 0
    for sid in sids:
        if emb_dim:
            fr, em = make_stream(sid, k, warm + steps, emb_dim=emb_dim)
            step = (lambda f: o.update(fr[f], em[f])) if tracker in ("botsort", "strongsort") else \
                (lambda f: o.update(fr[f], None, em[f]))
        else:
            fr = make_stream(sid, k, warm + steps)
            step = lambda f: o.update(fr[f])
        o = make()
        for f in range(warm):
            step(f)
        u0 = o.n_updates
        t0 = time.perf_counter()
        for f in range(warm, warm + steps):
            step(f)
        tot_t += time.perf_counter() - t0
        tot_u += o.n_updates - u0
    return tot_t, tot_u


def cpu_single_core(k, warm, steps, n_streams, tracker, emb_dim=0):
    t, u = _cpu_worker((list(range(n_streams)), k, warm, steps, tracker, emb_dim))
    return u / t, "%d streams x %d frames (after %d warm-up), K=%d, 1 core" % (n_streams, steps, warm, k)


def run_reference_arm(args):
    """--impl reference: the reference's CPU path (the oracle port: the reference is pure
    Python and /root/reference does not travel to the GPU box) on all host cores, one
    stream per worker process at a time, mirroring boxmot/engine/val.py:389-392."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import lap as olap
    olap.build()
    cores = os.cpu_count() or 1
    workers = max(1, cores - 1) if cores > 2 else cores
    per = max(1, args.ref_streams_per_core)
    steps = min(args.steps, 300)
    jobs = [([w * per + i for i in range(per)], args.objects, args.warmup, steps, args.tracker, args.emb_dim)
            for w in range(workers)]
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(workers) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    tmax = max(r[0] for r in res)
    upd = sum(r[1] for r in res)
    value = upd / tmax
    sample = "%d streams (%d per worker x %d workers) x %d frames after %d warm-up, K=%d" % (
        workers * per, per, workers, steps, args.warmup, args.objects)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "track-updates/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tmax / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s, synthetic streams x %d dets/frame (SURVEY 8(d) C2 streams), "
                               "CPU sample of the 1024-stream job" % (TRACKER_NAMES[args.tracker], args.objects),
                   "streams": workers * per, "objects": args.objects},
        "cpu_baseline": {"value": value, "unit": "track-updates/s", "cores": workers, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "track-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall, "host_cores": cores,
    }
    emit(line)


# --------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons every 100 ms during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz, self.ok = None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as ex:  # no NVML: report it, never invent numbers
            self.err = repr(ex)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [],
                    "note": "NVML unavailable or region shorter than one sample"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------- GPU arm
# DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) of ONE frame step at the default
# workload, from the ncu captures committed under profiles/ (cold caches: ncu flushes L2 before
# every kernel, so this is an upper bound of what a live step moves):
#   profiles/r1_launches_lane_scan_traffic.csv   predict 55.2 + assoc 19.3 + update 72.0 + finalize 39.7 MB
#   profiles/r1_launches_bytetrack_traffic.csv   byte_step_kernel 25.3 MB read + 15.8 MB written
#   profiles/r2_launches_ocsort_traffic.csv      round 2 (commit bf43990): predict 54.8 + assoc 19.3 + update 71.7 + finalize 39.5 MB
#   profiles/r2_launches_ocsort_traffic_final.csv  final kernels of round 2 (update at 128, finalize at 32 registers):
#                                                predict 53.1 + assoc 19.2 + update 72.3 + finalize 42.5 MB = 187.1 MB
NCU_TRAFFIC = {"ocsort": 187.1e6, "bytetrack": 41.1e6}
NCU_TRAFFIC_SOURCE = {"ocsort": "profiles/r2_launches_ocsort_traffic_final.csv (ncu dram__bytes_read+write of the four kernels of one "
                                "frame step, cold caches, single launch chain; captured with the final kernels of round 2)",
                      "bytetrack": "profiles/r1_launches_bytetrack_traffic.csv (round 1 capture)"}
TRACKER_NAMES = {"ocsort": "OC-SORT (ocsort.yaml)", "bytetrack": "ByteTrack (bytetrack.yaml)",
                 "botsort": "BoT-SORT (botsort.yaml, synthetic re-ID embeddings)",
                 "deepocsort": "DeepOCSORT (deepocsort.yaml, synthetic re-ID embeddings)",
                 "boosttrack": "BoostTrack++ (boosttrack.yaml, synthetic re-ID embeddings)",
                 "strongsort": "StrongSORT (constructor defaults, nn_budget 50, tracks born Confirmed; synthetic re-ID embeddings)"}


def alg_bytes(track_updates, matches, births, dets_in, rows_out):
    """Algorithmic bytes of one frame step (the SURVEY.md 8(d) terms, with the engine's
    block-form XYSR state of 7 + 13 fp64 = 160 B per track instead of the dense 7 + 49 =
    448 B the survey assumed): predict 2*160*T, update (2*160+32)*M, detections in 24*D
    (float32 x 6; the IoU operands), LAP result 4*(T+D), lifecycle scalars 64*T, output rows
    64*M_out, birth state write 160*B.  Cost matrix and LAP input count 0 bytes: they
    never leave the SM."""
    return (2 * 160 * track_updates + (2 * 160 + 32) * matches + 24 * dets_in
            + 4 * (track_updates + dets_in) + 64 * track_updates + 64 * rows_out + 160 * births)


def per_kernel(ksum_ms, steps, cnt, dets_in):
    """Live CUDA-event time of each of the four launches (averaged over `steps` L2-flushed
    steps) next to its share of the algorithmic bytes (the terms of alg_bytes)."""
    T, M, B = cnt["track_updates"], cnt["matches"], cnt["births"]
    nbytes = {"ocs_predict_kernel": 2 * 160 * T,
              "ocs_assoc_kernel": 24 * dets_in + 4 * (T + dets_in),
              "ocs_update_kernel": (2 * 160 + 32) * M,
              "ocs_finalize_kernel": 64 * T + 64 * M + 160 * B}
    tot = float(sum(ksum_ms)) or 1.0
    out = []
    for (name, nb), ms in zip(nbytes.items(), ksum_ms):
        out.append({"kernel": name, "avg_ms": float(ms) / steps, "share": float(ms) / tot,
                    "alg_bytes_per_launch": nb / steps,
                    "achieved_gbs": nb / (float(ms) * 1e-3) / 1e9 if ms > 0 else None})
    return out


def run_gpu_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    S, K, W = args.streams, args.steps, args.warmup
    F = W + K
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    gi = gen_inputs(rank * S, S, args.objects, F, args.det_cap, max(1, min(32, cores // world)), args.emb_dim)
    dets, nd = gi[0], gi[1]
    embs = gi[2] if args.emb_dim else None
    gen_s = time.perf_counter() - t0

    import torch
    import torch.distributed as dist
    from boxmot_b200.engine import (BoostTrackEngine, BotSortEngine, ByteTrackEngine, DeepOcSortEngine, OcSortEngine,
                                    StrongSortEngine)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    byte = args.tracker in ("bytetrack", "botsort")      # BYTE-family engines: no per-kernel timing API
    bot = args.tracker in EMB_TRACKERS       # embeddings travel with the detections
    deep, boost = args.tracker == "deepocsort", args.tracker == "boosttrack"
    strong = args.tracker == "strongsort"
    if strong:
        eng = StrongSortEngine(S, track_cap=args.track_cap, det_cap=args.det_cap, pool_cap=args.pool_cap, device=dev,
                               emb_dim=args.emb_dim, **STRONGSORT_ARGS)
        byte = False
    elif deep:
        eng = DeepOcSortEngine(S, track_cap=args.track_cap, det_cap=args.det_cap, device=dev, emb_dim=args.emb_dim,
                               **DEEPOCSORT_YAML)
        byte = False
    elif boost:
        eng = BoostTrackEngine(S, track_cap=args.track_cap, det_cap=args.det_cap, device=dev, emb_dim=args.emb_dim,
                               **BOOSTTRACK_YAML)
        byte = False
    elif bot:
        eng = BotSortEngine(S, track_cap=args.track_cap, det_cap=args.det_cap, device=dev, emb_dim=args.emb_dim,
                            **BOTSORT_YAML)
    elif byte:
        eng = ByteTrackEngine(S, track_cap=args.track_cap, det_cap=args.det_cap, device=dev, **BYTETRACK_YAML)
    else:
        eng = OcSortEngine(S, track_cap=args.track_cap, det_cap=args.det_cap, device=dev, **OCSORT_YAML)
    # OC-SORT batches of more than 1536 streams run the first-round screen as a fifth launch (capi.cu ocs_create)
    prescreen = (not (deep or boost or bot or byte)) and (
        os.environ.get("BM_PRESCREEN", "1" if S > 1536 else "0")[:1] == "1")
    n_kernels = 8 if strong else 7 if deep else 7 if boost else 3 if bot else 1 if byte else 5 if prescreen else 4
    groups = 1
    if hasattr(eng, "set_stream_groups"):
        groups = eng.set_stream_groups(args.groups if args.groups is not None else DEFAULT_GROUPS.get(args.tracker, 1))
    knames = KERNEL_NAMES.get(args.tracker)
    d_embs = None
    if bot:                              # device-resident embeddings, uploaded frame by frame (tens of GB)
        d_embs = torch.empty(embs.shape, dtype=torch.float32, device=dev)
        for f in range(F):
            d_embs[f].copy_(torch.from_numpy(embs[f]))
    h_dets = torch.from_numpy(dets).pin_memory()
    h_nd = torch.from_numpy(nd).pin_memory()
    d_dets = h_dets.to(dev, non_blocking=True)
    d_nd = h_nd.to(dev, non_blocking=True)
    flush_buf = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(step_fn):
        """W untimed steps, then exactly K steps, each bracketed by its own CUDA-event pair
        on the launching stream so the (optional) L2 flush between steps is not billed."""
        eng.reset()
        for f in range(W):
            step_fn(f)
        c0 = eng.counters()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        w0 = time.perf_counter()
        for i in range(K):
            if not args.no_flush:
                flush_buf.zero_()
            ev[i][0].record()
            step_fn(W + i)
            ev[i][1].record()
        barrier()
        wall = time.perf_counter() - w0
        sampler.stop_flag = True
        sampler.join()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        c1 = eng.counters()
        eng.status()   # raises on any capacity / solver error bit
        return ms, wall, {k: c1[k] - c0[k] for k in c1}, sampler.summary()

    # (1) inputs resident in HBM
    def step_dev(f):
        if bot:
            eng.step(d_dets[f], d_nd[f], d_embs[f])
        else:
            eng.step(d_dets[f], d_nd[f])

    ms_dev, wall_dev, cnt, clocks = timed(step_dev)

    # (1b) per-kernel breakdown of the same steps: the library brackets its four launches with
    # CUDA events on the launching stream (bm_ocsort_time_kernels); separate short pass so the
    # extra event records never sit inside the headline timing.
    KB = min(K, 100)
    ksum, kcnt = None, None
    if not byte:                       # the ByteTrack step is a single kernel: nothing to break down
        eng.reset()
        for f in range(W):
            step_dev(f)
        eng.time_kernels(True)
        ksum = np.zeros(len(knames))
        ck0 = eng.counters()
        for i in range(KB):
            if not args.no_flush:
                flush_buf.zero_()
            step_dev(W + i)
            ksum += eng.kernel_times()
        ck1 = eng.counters()
        eng.time_kernels(False)
        kcnt = {k: ck1[k] - ck0[k] for k in ck1}

    # (2) end to end through the public host-buffer API (boxmot_b200.engine.HostPipeline):
    # every step uploads that frame's detections from pinned host memory and downloads its
    # result rows; the copies of neighbouring frames overlap the association of the current
    # one.  Timed by the host clock between two full synchronisations (the device events of
    # one stream do not see the copy streams); no L2 flush here -- the inputs arrive from
    # host memory every step.
    e2e = None
    pipe = None
    if not args.no_e2e:
        from boxmot_b200.engine import HostPipeline
        eng.reset()
        pipe = HostPipeline(eng)
        rows_seen = 0
        # embeddings (when the tracker takes them): the K timed frames sit in pinned host memory before the clock
        # starts, as the detections do -- the timed region is H2D + step + D2H, not a host-side memcpy into a
        # staging buffer; the warm-up frames (untimed, tens of GB) go through two pinned staging buffers
        p_e = [torch.empty(embs.shape[1:], dtype=torch.float32).pin_memory() for _ in range(2)] if bot else None
        h_embs_timed = torch.from_numpy(embs[W:W + K]).pin_memory() if bot else None

        def submit(f):
            if bot:
                if f >= W:
                    return pipe.submit(h_dets[f], h_nd[f], h_embs_timed[f - W])
                p_e[f & 1].copy_(torch.from_numpy(embs[f]))
                return pipe.submit(h_dets[f], h_nd[f], p_e[f & 1])
            return pipe.submit(h_dets[f], h_nd[f])

        for f in range(W):
            submit(f)
        pipe.drain()
        c0 = eng.counters()
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            r = submit(W + i)
            if r is not None:
                rows_seen += len(r[0])                # the host consumes every frame's result
                pipe.account(r)
        r = pipe.drain()
        rows_seen += len(r[0])
        pipe.account(r)
        barrier()
        ms_e2e = (time.perf_counter() - t0) * 1e3
        c1 = eng.counters()
        eng.status()
        e2e = (ms_e2e, c1["track_updates"] - c0["track_updates"])

    # aggregate over ranks: max time, summed units
    def agg(ms, units):
        if world == 1:
            return ms, units
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        u = torch.tensor([units], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(u, op=dist.ReduceOp.SUM)
        return float(t.item()), float(u.item())

    ms_all, upd_all = agg(ms_dev, cnt["track_updates"])
    line = None
    if e2e is not None:
        ms_e_all, upd_e_all = agg(*e2e)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import lap as olap
        olap.build()
        v, sample = cpu_single_core(args.objects, min(W, 10), min(K, 290), 1 if args.tracker == "botsort" else
                                    2 if bot else args.cpu_streams,
                                    args.tracker, args.emb_dim)
        cpu = {"value": v, "unit": "track-updates/s", "cores": 1, "kind": "port", "sample": sample}
    if rank == 0:
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            peaks = json.load(open(pk))
        peak = peaks.get("hbm_gbs", 6650.0)
        dets_in = int(nd[W:W + K].sum())
        # output rows <= matched tracks (rows are emitted only for tracks updated this frame)
        nbytes = alg_bytes(cnt["track_updates"], cnt["matches"], cnt["births"], dets_in, cnt["matches"])
        if bot:
            # + embeddings: detection features in (4E per detection), smooth features read by the
            # contraction (4E per live track ~ track-updates), EMA read+write (2*4E per match),
            # distance matrix out and in again (2 * 4 bytes per track x detection pair)
            E = args.emb_dim
            nbytes += 4 * E * dets_in + 4 * E * cnt["track_updates"] + 8 * E * cnt["matches"] \
                + 8 * cnt["track_updates"] * (dets_in // max(1, K * S))
            if strong:
                # the sample bank: every distinct stored vector is read once by the contraction (fp32 shadow,
                # 4E bytes; the reference's nominal T*B*E is ~10x that because its lists hold each vector ~10
                # times); fp64 cost matrix out and in (16 bytes per pair); new feature fp64 + fp32 per match
                nbytes += 4 * E * cnt.get("bank_rows", 0) + 16 * cnt["track_updates"] * (dets_in // max(1, K * S)) \
                    + 12 * E * cnt["matches"]
            if deep or boost:
                # the first-round matrix is materialised (column weights / column softmax): fp64 out and in
                nbytes += 16 * cnt["track_updates"] * (dets_in // max(1, K * S))
        achieved = nbytes / (ms_dev * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": upd_all / (ms_all * 1e-3), "unit": "track-updates/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_all / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": ("DeepOCSORT (deepocsort.yaml) on %d concurrent synthetic streams/GPU x %d objects/frame x %d frames, "
                                    "512-d re-ID embeddings: tcgen05 embedding contraction + adaptive top-2 weighting + IoU/OCM cost, "
                                    "lapjv-exact assignment, embedding EMA (tracker of BASELINE configs[4] on the configs[1] streams)" if deep else
                                    "StrongSORT (constructor defaults, nn_budget 50, occlusion handler off, tracks born Confirmed = the fork's "
                                    "runnable mode) on %d concurrent synthetic streams/GPU x %d objects/frame x %d frames, 512-d re-ID "
                                    "embeddings: tcgen05 sample-bank contraction, Mahalanobis-gated three-stage cascade with scipy-exact "
                                    "assignments, NSA Kalman update, adaptive feature EMA, id recovery, partial_fit (BASELINE configs[3])" if strong else
                                    "BoostTrack++ (boosttrack.yaml) on %d concurrent synthetic streams/GPU x %d objects/frame x %d frames, "
                                    "512-d re-ID embeddings: DLO/DUO boosts, soft-BIoU / Mahalanobis / shape / embedding score, "
                                    "lapjv-exact assignment (tracker of BASELINE configs[4] on the configs[1] streams)" if boost else
                                    "BoT-SORT (botsort.yaml) on %d concurrent synthetic streams/GPU x %d objects/frame x %d frames, "
                                    "512-d re-ID embeddings: cosine-distance contraction + IoU fusion + Kalman XYWH "
                                    "(BASELINE configs[2])" if bot else
                                    "ByteTrack (bytetrack.yaml) on %d concurrent synthetic streams/GPU x %d objects/frame x %d frames, "
                                    "IoU + fuse_score cost, three thresholded lapjv-exact assignments per frame "
                                    "(tracker of BASELINE configs[0] on the configs[1] streams)" if byte else
                                    "OC-SORT (ocsort.yaml) on %d concurrent synthetic streams/GPU x %d objects/frame x %d frames, "
                                    "IoU + OCM velocity cost, lapjv-exact assignment (BASELINE configs[1])").replace("512-d", "%d-d" % args.emb_dim) % (S, args.objects, F),
                       "tracker": args.tracker,
                       "emb_dim": args.emb_dim if args.tracker in EMB_TRACKERS else 0,
                       "stream_groups": groups,
                       "streams_per_gpu": S, "objects": args.objects, "frames": F, "det_cap": args.det_cap,
                       "track_cap": args.track_cap, "partition": "streams sharded by rank, no data-path collective",
                       "l2": "no flush" if args.no_flush else "L2 flushed (252 MB memset) between timed steps, outside the event pairs"},
            "stream_frames_per_s": world * S * K / (ms_all * 1e-3),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC.get(args.tracker) if (S, args.objects, args.det_cap, args.track_cap) ==
                         (1024, 100, 128, 256) else None,
                         "traffic_source": NCU_TRAFFIC_SOURCE.get(args.tracker, "not captured for this configuration"),
                         "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback",
                         "kernel": "frame step = " + " + ".join(knames) if (deep or boost or strong) else
                                   "frame step = bot_embed + byte_step<BoT> + bot_ema" if bot else
                                   "byte_step_kernel" if byte else
                                   "frame step = ocs_predict + %socs_assoc + ocs_update + ocs_finalize" % ("ocs_screen + " if prescreen else ""),
                         "alg_bytes_per_launch": nbytes / K, "avg_launch_ms": ms_dev / K},
            "gpu_launches": n_kernels * groups * K * (2 if not args.no_e2e else 1) + (0 if args.no_e2e else K),
            "clocks": clocks,
            "counters": cnt, "input_gen_s": gen_s, "wall_s_timed_loop": wall_dev,
        }
        if ksum is not None and not (deep or boost or strong):
            line["roofline"]["per_kernel"] = per_kernel(ksum, KB, kcnt, int(nd[W:W + KB].sum()))
            line["roofline"]["dominant_kernel"] = "ocs_assoc_kernel"
        elif ksum is not None:
            tot = float(ksum.sum()) or 1.0
            line["roofline"]["per_kernel"] = [{"kernel": n, "avg_ms": float(m) / KB, "share": float(m) / tot}
                                              for n, m in zip(knames, ksum)]
            line["roofline"]["dominant_kernel"] = knames[int(np.argmax(ksum))]
        if e2e is not None:
            line["e2e"] = {"value": upd_e_all / (ms_e_all * 1e-3), "unit": "track-updates/s",
                           "h2d_bytes_per_step": int(pipe.h2d_bytes), "d2h_bytes_per_step": int(pipe.d2h_bytes),
                           "timing": "host clock over K pipelined steps between full syncs, no L2 flush; rows are packed on the device and written "
                                     "by the kernel into mapped pinned host memory (64 B per emitted row + offsets)",
                           "ms_per_step": ms_e_all / K}
        if cpu is not None:
            line["cpu_baseline"] = cpu
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    # free the device before a sub-run of another configuration
    eng.close()
    del eng, d_dets, d_nd, d_embs, h_dets, h_nd, flush_buf, pipe
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return line


# --------------------------------------------------------------------------- the other BASELINE configs
# Short driver-timed sub-runs of the configurations the headline line does not cover (single GPU only),
# attached to the headline line as "configs": device value, ms/step, roofline fraction, e2e.
SUB_CONFIGS = [
    ("C3_botsort_1024x200_e512", ["--tracker", "botsort", "--steps", "10", "--warmup", "10"]),
    ("C4_strongsort_512x100_e768_b50", ["--tracker", "strongsort", "--steps", "8", "--warmup", "70"]),
    ("C5_deepocsort_crowd_1024x480_e512", ["--tracker", "deepocsort", "--objects", "480", "--det-cap", "512",
                                          "--track-cap", "768", "--emb-dim", "512", "--steps", "4", "--warmup", "6"]),
    ("C5_boosttrack_crowd_1024x480_e512", ["--tracker", "boosttrack", "--objects", "480", "--det-cap", "512",
                                           "--track-cap", "768", "--emb-dim", "512", "--steps", "4", "--warmup", "6"]),
]


def c1_single_stream_latency():
    """BASELINE configs[0]: create_tracker("bytetrack").update() on MOT17-mini det.txt (MOT17-04-FRCNN, the
    fixture tests/golden/mot17_mini_dets.npz), one stream, frame by frame from host arrays: the latency the
    reference's own performance test bounds (tests/performance/test_tracking_p.py:53-58, < 5 ms), next to the
    oracle port on one host core."""
    import boxmot_b200
    from oracle.bytetrack import ByteTrackOracle
    z = np.load(os.path.join(ROOT, "tests", "golden", "mot17_mini_dets.npz"))["MOT17_04_FRCNN"]
    n = int(z[:, 0].max())
    frames = [np.c_[z[z[:, 0] == f][:, 1:], np.zeros((z[:, 0] == f).sum())].astype(np.float64) for f in range(1, n + 1)]
    img = np.zeros((1080, 1920, 3), np.uint8)
    trk = boxmot_b200.create_tracker("bytetrack", boxmot_b200.get_tracker_config("bytetrack"))
    for d in frames[:50]:
        trk.update(d, img)
    t0 = time.perf_counter()
    for d in frames[50:]:
        trk.update(d, img)
    gpu_ms = (time.perf_counter() - t0) * 1e3 / max(1, len(frames) - 50)
    o = ByteTrackOracle(**BYTETRACK_YAML)
    for d in frames[:50]:
        o.update(d)
    t0 = time.perf_counter()
    for d in frames[50:]:
        o.update(d)
    cpu_ms = (time.perf_counter() - t0) * 1e3 / max(1, len(frames) - 50)
    return {"workload": "ByteTrack on MOT17-mini MOT17-04-FRCNN det.txt, single stream, update() from host arrays",
            "frames": len(frames) - 50, "update_ms": gpu_ms, "oracle_port_update_ms": cpu_ms,
            "note": "one stream cannot fill a GPU: this is launch + copy latency, not throughput"}


def run_sub_configs(headline_argv):
    import copy
    out = {}
    for name, extra in SUB_CONFIGS:
        argv = [a for a in headline_argv if a in ("--no-flush",)] + extra + ["--no-cpu-baseline"]
        old = sys.argv
        try:
            sys.argv = [old[0]] + argv
            a = parse()
            t0 = time.perf_counter()
            line = run_gpu_arm(a)
            out[name] = {"value": line["value"], "unit": line["unit"], "ms_per_step": line["ms_per_step"],
                         "steps": line["steps"], "warmup": line["warmup"],
                         "roofline_frac": line["roofline"]["frac"], "achieved_gbs": line["roofline"]["achieved"],
                         "dominant_kernel": line["roofline"].get("dominant_kernel"),
                         "per_kernel_ms": {k["kernel"]: k["avg_ms"] for k in line["roofline"].get("per_kernel", [])},
                         "e2e": line.get("e2e"), "config": line["config"], "counters": line["counters"],
                         "wall_s": time.perf_counter() - t0}
        except Exception as ex:        # a sub-run must never take the headline line down
            out[name] = {"error": repr(ex)[:300]}
        finally:
            sys.argv = old
    try:
        out["C1_bytetrack_mot17mini_single_stream"] = c1_single_stream_latency()
    except Exception as ex:
        out["C1_bytetrack_mot17mini_single_stream"] = {"error": repr(ex)[:300]}
    return out


def main():
    args = parse()
    # stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version
    # banner, for one) is diverted to stderr by pointing fd 1 at fd 2 while the run lasts.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
        return
    line = run_gpu_arm(args)
    if line is not None and int(os.environ.get("WORLD_SIZE", "1")) == 1 and args.tracker == "ocsort" \
            and args.sub_configs and (args.streams, args.objects) == (1024, 100):
        line["configs"] = run_sub_configs(sys.argv[1:])
    if line is not None:
        emit(line)


if __name__ == "__main__":
    main()
