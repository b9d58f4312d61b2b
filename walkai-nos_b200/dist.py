"""Multi-GPU plumbing for the scorer: one process per GPU, nodes sharded contiguously
across ranks, pods replicated.  Scoring needs no data-path collective; the only exchange is the
per-demand "first node that would update" key, merged with one all-reduce(MIN) per pass — first node
wins across shards exactly as tryRepartition does over the whole list (mig_controller.go:146-166).
Works on CPU tensors with the gloo backend (tests) and CUDA tensors with NCCL (bench)."""
from __future__ import annotations

import numpy as np

NO_NODE = 0x7FFFFFFF


def shard_bounds(n_nodes: int, rank: int, world: int):
    """Rank r owns nodes [r*N/world, (r+1)*N/world): contiguous, so the lowest global index wins."""
    return rank * n_nodes // world, (rank + 1) * n_nodes // world


def shard_snapshot(snap, rank: int, world: int):
    """The rank's slice of a flattened snapshot (dict of arrays) and its first global node index."""
    if snap is None:
        return None, 0
    N = int(snap["n_nodes"])
    lo, hi = shard_bounds(N, rank, world)
    g0, g1 = int(snap["gpu_off"][lo]), int(snap["gpu_off"][hi])
    out = dict(n_nodes=hi - lo, gpu_off=(snap["gpu_off"][lo:hi + 1] - g0).astype(np.int32), valid=snap["valid"][lo:hi])
    for k in ("model", "mem_gb", "used", "free", "used_mask", "free_mask"):
        if k in snap:
            out[k] = snap[k][g0:g1]
    return out, lo


def merge_first_node(first_local, lo: int, group=None):
    """first_local: int32 torch tensor [D] of LOCAL node indices (-1 = none) on any device.
    Returns the GLOBAL winner per demand (-1 = none) after one all-reduce(MIN)."""
    import torch
    import torch.distributed as dist
    key = torch.where(first_local >= 0, first_local + lo, torch.full_like(first_local, NO_NODE))
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(key, op=dist.ReduceOp.MIN, group=group)
    return torch.where(key == NO_NODE, torch.full_like(key, -1), key)


def merge_first_node_keys(first_local, lo_vec, group=None):
    """Same as merge_first_node for several demand sets at once: first_local [D_total] holds local indices
    of different node lists back to back and lo_vec [D_total] the matching shard offsets, so one
    all-reduce(MIN) resolves all of them (collectives here are latency-bound: issue as few as possible)."""
    import torch
    import torch.distributed as dist
    key = torch.where(first_local >= 0, first_local + lo_vec, NO_NODE)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(key, op=dist.ReduceOp.MIN, group=group)
    return torch.where(key == NO_NODE, -1, key)


def open_exchange(planner, max_keys: int, group=None):
    """Sets up the library's own first-node exchange (peer-memory atomicMin over NVLink, include/nosplan.h
    "sharded scoring"): every rank creates its buffer, the 64-byte IPC handles travel through the
    process group once, every rank maps its peers.  After this, planner.exchange_min_dev replaces
    merge_first_node_keys' all-reduce.  Raises NosplanError when peer mapping is not possible."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    rank = dist.get_rank(group) if world > 1 else 0
    handle = planner.exchange_create(rank, world, max_keys)
    handles = [handle]
    if world > 1:
        handles = [None] * world
        dist.all_gather_object(handles, handle, group=group)
    planner.exchange_open(handles)
    if world > 1:
        dist.barrier(group=group)  # nobody pushes before everybody has mapped


def close_exchange(planner, group=None):
    """Tears the exchange down safely: every rank first confirms its last pass has finished (status call), then the
    ranks meet at a barrier, and only then is any buffer freed — a peer still in flight would otherwise write its step
    flag into freed memory (include/nosplan.h, "Rules")."""
    import torch.distributed as dist
    try:
        planner.exchange_status()
    finally:
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.barrier(group=group)
        planner.exchange_destroy()


def merge_presence(count_local, group=None):
    """count_local: int64 torch tensor [n_keys], this shard's sum of used + free per key (Planner.presence).
    Returns the cluster-wide sums after one all-reduce(SUM): Controller.profileAlreadyPresent
    (mig_controller.go:121-144) holds for a pod iff every key it requests has a positive sum."""
    import torch.distributed as dist
    out = count_local.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out
