"""Stream sharding across GPUs.

Streams are independent trackers -- the reference runs one tracker per stream
(boxmot/engine/track.py:41-53) or one process per sequence (boxmot/engine/val.py:389-392)
-- so the path partitions with NO collective inside the frame loop: rank r owns a
contiguous block of streams and its own engine handle.  The only exchange is the optional
gather of output rows (and of the scalar timings in bench.py) to rank 0, at most once per
frame and normally once per run.  Works with the ``nccl`` backend on GPUs and ``gloo`` on CPU
(the tests run world_size 2 over gloo)."""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(n_streams, rank, world):
    """Contiguous block [lo, hi) of the global streams owned by ``rank``; the first
    ``n_streams % world`` ranks take one extra stream."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, extra = divmod(int(n_streams), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def owner_of(stream, n_streams, world):
    """Rank that owns global stream index ``stream`` (inverse of ``shard_range``)."""
    base, extra = divmod(int(n_streams), int(world))
    cut = extra * (base + 1)
    if stream < cut:
        return stream // (base + 1)
    return extra + (stream - cut) // base


def compact_rows(out, nout):
    """[S, R, 8] padded rows + [S] counts -> ([sum(nout), 9] rows prefixed by the LOCAL
    stream index, counts) -- the ragged form that is exchanged."""
    out = np.asarray(out)
    nout = np.asarray(nout).astype(np.int64)
    S, R = out.shape[0], out.shape[1]
    mask = np.arange(R)[None, :] < nout[:, None]
    sid = np.broadcast_to(np.arange(S, dtype=np.float64)[:, None], (S, R))[mask]
    return np.concatenate([sid[:, None], out[mask]], axis=1), nout


def gather_rows(out, nout, n_streams, group=None, device=None):
    """Gather every rank's output rows of one frame on rank 0.

    ``out`` [S_local, R, 8] float64 and ``nout`` [S_local] int32 (numpy or tensors) are this
    rank's engine outputs for its ``shard_range``.  Returns on rank 0 a list of
    ``n_streams`` arrays (M_s, 8) in GLOBAL stream order, on other ranks ``None``.
    Two collectives: an all_gather of the per-rank row totals, then an all_gather of the
    rows padded to the largest total (NVSwitch makes the padded form cheaper than a
    send/recv fan-in at these sizes: <= 8 MB per rank per frame)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if torch.is_tensor(out):
        out, nout = out.cpu().numpy(), nout.cpu().numpy()
    lo, hi = shard_range(n_streams, rank, world)
    if out.shape[0] != hi - lo:
        raise ValueError("rank %d holds %d streams, its shard has %d" % (rank, out.shape[0], hi - lo))
    rows, _ = compact_rows(out, nout)
    rows[:, 0] += lo                                    # global stream index
    dev = torch.device(device) if device is not None else (
        torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl"
        else torch.device("cpu"))
    tot = torch.tensor([rows.shape[0]], dtype=torch.int64, device=dev)
    tots = [torch.zeros_like(tot) for _ in range(world)]
    dist.all_gather(tots, tot, group=group)
    tots = [int(t.item()) for t in tots]
    pad = max(max(tots), 1)
    buf = torch.zeros((pad, 9), dtype=torch.float64, device=dev)
    if rows.shape[0]:
        buf[: rows.shape[0]] = torch.from_numpy(rows).to(dev)
    bufs = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf, group=group)
    if rank != 0:
        return None
    allrows = np.concatenate([b[:n].cpu().numpy() for b, n in zip(bufs, tots)], axis=0)
    sid = allrows[:, 0].astype(np.int64)
    # rows of one stream are contiguous and in the reference's order already
    bounds = np.searchsorted(sid, np.arange(n_streams + 1))
    return [allrows[bounds[s]:bounds[s + 1], 1:] for s in range(n_streams)]


def reduce_timing(ms_local, units_local, group=None, device=None):
    """bench.py's aggregation: (max over ranks of the device time, sum of the units)."""
    dev = torch.device(device) if device is not None else (
        torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl"
        else torch.device("cpu"))
    t = torch.tensor([float(ms_local)], dtype=torch.float64, device=dev)
    u = torch.tensor([float(units_local)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(u, op=dist.ReduceOp.SUM, group=group)
    return float(t.item()), float(u.item())
