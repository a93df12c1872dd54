"""Deterministic synthetic detection / embedding streams.

Per stream ``s`` the generator is seeded with ``default_rng(1234 + s)``: K persistent
objects in a 1920x1080 frame with constant velocity and reflection at the borders, each
emitting a detection with probability 0.95 (box jitter N(0, 2 px) per coordinate,
conf ~ U(0.05, 0.99)), plus Poisson(0.05 K) false positives; detection order is
shuffled per frame.  Embeddings are unit vectors id_k + 0.05 N(0, I), re-normalised.

This is the *input* side of the hot path (what the reference gets from a detector and
a ReID model: boxmot/utils/dataloaders/MOT17.py:147-200 slices the same arrays from
disk), not part of the association arithmetic.
"""
import numpy as np

W, H = 1920, 1080


def make_stream(stream_id, n_objects, n_frames, emb_dim=0, crowd=False, seed_base=1234):
    """Return a list of per-frame ``dets`` (N,6) float64 arrays (and ``embs`` (N,E)
    float32 arrays when ``emb_dim`` > 0) for one stream."""
    rng = np.random.default_rng(seed_base + stream_id)
    K = n_objects
    if crowd:
        # MOT20-like density: centres clustered around a few hot spots.
        n_hot = max(1, K // 50)
        hot = np.c_[rng.uniform(300, W - 300, n_hot), rng.uniform(250, H - 250, n_hot)]
        pick = rng.integers(0, n_hot, K)
        cx = hot[pick, 0] + rng.normal(0, 140, K)
        cy = hot[pick, 1] + rng.normal(0, 110, K)
        cx = np.clip(cx, 100, W - 100)
        cy = np.clip(cy, 100, H - 100)
    else:
        cx = rng.uniform(100, W - 100, K)
        cy = rng.uniform(100, H - 100, K)
    bw = rng.uniform(30, 120, K)
    bh = bw * rng.uniform(1.5, 3.0, K)
    vx = rng.normal(0, 3, K)
    vy = rng.normal(0, 3, K)
    ident = None
    if emb_dim:
        ident = rng.normal(0, 1, (K, emb_dim))
        ident /= np.linalg.norm(ident, axis=1, keepdims=True)
    frames, embs = [], []
    for _ in range(n_frames):
        cx = cx + vx
        cy = cy + vy
        bounce = (cx < 100) | (cx > W - 100)
        vx = np.where(bounce, -vx, vx)
        cx = np.clip(cx, 100, W - 100)
        bounce = (cy < 100) | (cy > H - 100)
        vy = np.where(bounce, -vy, vy)
        cy = np.clip(cy, 100, H - 100)
        seen = rng.random(K) < 0.95
        idx = np.nonzero(seen)[0]
        n_fp = rng.poisson(0.05 * K)
        n = idx.size + n_fp
        boxes = np.empty((n, 4))
        boxes[: idx.size, 0] = cx[idx] - bw[idx] / 2
        boxes[: idx.size, 1] = cy[idx] - bh[idx] / 2
        boxes[: idx.size, 2] = cx[idx] + bw[idx] / 2
        boxes[: idx.size, 3] = cy[idx] + bh[idx] / 2
        boxes[: idx.size] += rng.normal(0, 2, (idx.size, 4))
        fcx = rng.uniform(100, W - 100, n_fp)
        fcy = rng.uniform(100, H - 100, n_fp)
        fw = rng.uniform(30, 120, n_fp)
        fh = fw * rng.uniform(1.5, 3.0, n_fp)
        boxes[idx.size:, 0] = fcx - fw / 2
        boxes[idx.size:, 1] = fcy - fh / 2
        boxes[idx.size:, 2] = fcx + fw / 2
        boxes[idx.size:, 3] = fcy + fh / 2
        conf = rng.uniform(0.05, 0.99, n)
        d = np.c_[boxes, conf, np.zeros(n)]
        e = None
        if emb_dim:
            e = np.empty((n, emb_dim))
            e[: idx.size] = ident[idx] + 0.05 * rng.normal(0, 1, (idx.size, emb_dim))
            e[idx.size:] = rng.normal(0, 1, (n_fp, emb_dim))
            e /= np.linalg.norm(e, axis=1, keepdims=True)
            e = e.astype(np.float32)
        perm = rng.permutation(n)
        frames.append(d[perm])
        if emb_dim:
            embs.append(e[perm])
    return (frames, embs) if emb_dim else frames


def make_batch(n_streams, n_objects, n_frames, d_max, emb_dim=0, crowd=False,
               first_stream=0, seed_base=1234):
    """Padded batch: ``dets[F, S, d_max, 6]`` float64, ``ndets[F, S]`` int32
    (and ``embs[F, S, d_max, E]`` float32).  Frames with more than ``d_max``
    detections are an error (the engine does not silently truncate)."""
    dets = np.zeros((n_frames, n_streams, d_max, 6))
    nd = np.zeros((n_frames, n_streams), dtype=np.int32)
    embs = np.zeros((n_frames, n_streams, d_max, emb_dim), dtype=np.float32) if emb_dim else None
    for s in range(n_streams):
        out = make_stream(first_stream + s, n_objects, n_frames, emb_dim, crowd, seed_base)
        fr, em = out if emb_dim else (out, None)
        for f, d in enumerate(fr):
            if len(d) > d_max:
                raise ValueError("frame has %d detections > d_max=%d" % (len(d), d_max))
            dets[f, s, : len(d)] = d
            nd[f, s] = len(d)
            if emb_dim:
                embs[f, s, : len(d)] = em[f]
    return (dets, nd, embs) if emb_dim else (dets, nd)


def mot17_embs(frames, emb_dim, seed):
    """Seeded unit-norm float32 embeddings for det.txt rows (det.txt carries no identities, so the
    embedding is a smooth function of the box centre + noise: neighbouring boxes of consecutive
    frames get similar vectors, which is what makes the embedding term matter)."""
    rng = np.random.default_rng(seed)
    proj = rng.normal(0, 1, (4, emb_dim))
    out = []
    for d in frames:
        c = np.c_[(d[:, 0] + d[:, 2]) / 400.0, (d[:, 1] + d[:, 3]) / 400.0, (d[:, 2] - d[:, 0]) / 100.0,
                  np.ones(len(d))]
        e = np.sin(c @ proj) + 0.05 * rng.normal(0, 1, (len(d), emb_dim))
        e /= np.linalg.norm(e, axis=1, keepdims=True) + 1e-12
        out.append(e.astype(np.float32))
    return out


def strongsort_stream(stream_id, n_objects, n_frames, emb_dim, crowd=False):
    """make_stream for the StrongSORT fixtures: detections rounded to float32 (the engine's input
    precision) and returned as float64, embeddings float32 values as float64 (so every operation of
    the reference is float64).  Stream 65 also gets frames without detections, a frame with nothing
    above min_conf and a gap long enough to delete tracks and recover them from the lost buffer."""
    fr, em = make_stream(stream_id, n_objects, n_frames, emb_dim=emb_dim, crowd=crowd)
    fr = [d.astype(np.float32).astype(np.float64) for d in fr]
    em = [x.astype(np.float64) for x in em]
    if stream_id == 65:
        fr[7] = np.empty((0, 6))
        em[7] = np.empty((0, emb_dim))
        fr[20][:, 4] = 0.05
        for i in range(40, 49):
            fr[i] = fr[i][:3]
            em[i] = em[i][:3]
    return fr, em
