"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL on the GPU box, gloo in CPU tests).

The path shards with NO data-path collective: candidates split into contiguous global
index ranges, restarts split round-robin; training data and the factor are replicated.  The only
exchanges are
  * one all-gather of (value f64, index i64) per acquisition call -- NCCL has no MAXLOC, so each
    rank gathers the 16-byte pairs and reduces locally with the same lowest-index tie-break the
    kernel uses, which makes the result independent of the number of ranks;
  * one all-gather of (-LML, theta[p]) per fit to pick the best restart.
"""
import math

import numpy as np


def _dist():
    import torch.distributed as td
    return td


def rank_world(group=None):
    """(rank, world) of ``group``; (0, 1) when ``group`` is None (single process, no collective)."""
    if group is None:
        return 0, 1
    td = _dist()
    g = None if group == 'world' else group
    return td.get_rank(g), td.get_world_size(g)


def shard_range(m, rank, world):
    """contiguous global index range [lo, hi) of ``rank``: ranges tile [0, m) in rank order"""
    base, rem = divmod(int(m), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _comm_device(group):
    import torch
    td = _dist()
    g = None if group == 'world' else group
    return torch.device('cuda', torch.cuda.current_device()) if td.get_backend(g) == 'nccl' else torch.device('cpu')


def reduce_argmax_pairs(values, indices):
    """pure reduction rule: highest value wins, ties -> lowest index, NaN / index<0 never win."""
    best_v, best_i = -math.inf, -1
    for v, i in zip(values, indices):
        v, i = float(v), int(i)
        if i < 0 or math.isnan(v):
            continue
        if best_i < 0 or v > best_v or (v == best_v and i < best_i):
            best_v, best_i = v, i
    return best_v, best_i


def allreduce_argmax(value, index, group=None):
    """global (value, index) over the ranks of ``group``; identity when group is None"""
    if group is None:
        return value, index
    import torch
    td = _dist()
    g = None if group == 'world' else group
    world = td.get_world_size(g)
    dev = _comm_device(group)
    # bit-cast the f64 value into an i64 lane so ONE all-gather carries the 16-byte pair
    mine = torch.tensor([np.float64(value).view(np.int64).item(), int(index)], dtype=torch.int64, device=dev)
    out = torch.empty(2 * world, dtype=torch.int64, device=dev)
    td.all_gather_into_tensor(out, mine, group=g)
    pairs = out.cpu().numpy().reshape(world, 2)
    vals = pairs[:, 0].copy().view(np.float64)
    return reduce_argmax_pairs(vals, pairs[:, 1])


def merge_top(values, indices, k):
    """pure rule: the k best distinct (value, index) pairs, best first, ties -> lowest index"""
    values, indices = np.asarray(values, dtype=np.float64), np.asarray(indices, dtype=np.int64)
    keep = (indices >= 0) & ~np.isnan(values)
    values, indices = values[keep], indices[keep]
    indices, first = np.unique(indices, return_index=True)
    values = values[first]
    order = np.lexsort((indices, -values))[:k]
    return values[order], indices[order]


def allgather_top(values, indices, k, group=None):
    """every rank contributes its local top candidates; every rank returns the same global top-k"""
    values, indices = merge_top(values, indices, k)
    if group is None:
        return values, indices
    import torch
    td = _dist()
    g = None if group == 'world' else group
    world = td.get_world_size(g)
    dev = _comm_device(group)
    mine = np.full((k, 2), -1, dtype=np.int64)
    mine[:len(indices), 0] = values.view(np.int64)
    mine[:len(indices), 1] = indices
    out = torch.empty(2 * k * world, dtype=torch.int64, device=dev)
    td.all_gather_into_tensor(out, torch.from_numpy(mine.reshape(-1)).to(dev), group=g)
    pairs = out.cpu().numpy().reshape(world * k, 2)
    return merge_top(pairs[:, 0].copy().view(np.float64), pairs[:, 1], k)


def select_best_restart(thetas, neg_lmls, p, group):
    """each rank contributes its best local restart; every rank returns the global (theta, -lml).
    Ties -> lowest rank.  Raises LinAlgError if no rank has a finite restart."""
    import torch
    td = _dist()
    g = None if group == 'world' else group
    world = td.get_world_size(g)
    dev = _comm_device(group)
    row = np.full(p + 1, np.inf)
    if len(neg_lmls) and np.isfinite(neg_lmls).any():
        b = int(np.argmin(neg_lmls))
        row[0], row[1:] = neg_lmls[b], thetas[b]
    mine = torch.from_numpy(row).to(dev)
    out = torch.empty(world * (p + 1), dtype=torch.float64, device=dev)
    td.all_gather_into_tensor(out, mine, group=g)
    table = out.cpu().numpy().reshape(world, p + 1)
    if not np.isfinite(table[:, 0]).any():
        raise np.linalg.LinAlgError("no restart produced a positive-definite kernel matrix")
    b = int(np.argmin(table[:, 0]))
    return table[b, 1:].copy(), float(table[b, 0])


def broadcast_array(a, group=None, src=0):
    """the float64 array of rank ``src`` on every rank (shape included); identity when group is None"""
    if group is None:
        return a
    import torch
    td = _dist()
    g = None if group == 'world' else group
    dev = _comm_device(group)
    src_global = td.get_global_rank(g, src) if g is not None else src
    a = np.ascontiguousarray(a, dtype=np.float64)
    shape = torch.tensor(list(a.shape) + [-1] * (8 - a.ndim), dtype=torch.int64, device=dev)
    td.broadcast(shape, src=src_global, group=g)
    dims = [int(v) for v in shape.cpu().tolist() if v >= 0]
    buf = torch.from_numpy(a).to(dev) if td.get_rank(g) == src else torch.empty(dims, dtype=torch.float64, device=dev)
    td.broadcast(buf, src=src_global, group=g)
    return buf.cpu().numpy()
