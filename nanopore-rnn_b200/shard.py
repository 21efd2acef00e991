"""Read-level sharding across GPUs: windows are independent units and BN
groups never span reads, so reads are assigned whole to ranks by longest-processing-time
greedy on their window counts; no collective on the data path, labels are gathered on the host
(or with one all_gather_object when a single process needs them all)."""
import heapq

import numpy as np


def assign_reads(window_counts, world_size):
    """LPT greedy: returns ``world_size`` lists of read indices (each sorted ascending so a rank
    keeps its reads in input order).  Deterministic: ties break on the lower read index / rank."""
    counts = np.asarray(window_counts, dtype=np.int64)
    order = sorted(range(len(counts)), key=lambda i: (-int(counts[i]), i))
    heap = [(0, r) for r in range(world_size)]
    heapq.heapify(heap)
    out = [[] for _ in range(world_size)]
    for i in order:
        load, r = heapq.heappop(heap)
        out[r].append(i)
        heapq.heappush(heap, (load + int(counts[i]), r))
    return [sorted(x) for x in out]


def shard_loads(window_counts, assignment):
    counts = np.asarray(window_counts, dtype=np.int64)
    return [int(counts[a].sum()) if len(a) else 0 for a in assignment]


def gather_labels(local, rank, world_size, group=None):
    """Host gather of per-read results: ``local`` = {read index: value}.  With torch.distributed
    initialised, returns the merged dict on every rank; single process: returns ``local``."""
    if world_size == 1:
        return dict(local)
    import torch.distributed as dist
    bucket = [None] * world_size
    dist.all_gather_object(bucket, local, group=group)
    merged = {}
    for part in bucket:
        merged.update(part)
    return merged


def length_sorted_order(window_counts):
    """Reads in descending window count (stable): the packing order that keeps short tails together."""
    counts = np.asarray(window_counts, dtype=np.int64)
    return np.argsort(-counts, kind="stable")
