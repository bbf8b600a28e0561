"""Multi-GPU sharding of a document batch: documents are independent, so a batch
is partitioned across ranks by input bytes (greedy longest-processing-time bin packing), each
rank parses its shard on its own GPU, and the results are gathered on the host in the original
order.  No data-path collective: torch.distributed is only the host-side gather."""
from __future__ import annotations

import heapq
from typing import Callable, Sequence


def partition(sizes: Sequence[int], n_shards: int) -> "list[list[int]]":
    """LPT bin packing: indices per shard, deterministic, each shard's indices ascending."""
    heap = [(0, s) for s in range(n_shards)]
    heapq.heapify(heap)
    shards = [[] for _ in range(n_shards)]
    for i in sorted(range(len(sizes)), key=lambda i: (-sizes[i], i)):
        load, s = heapq.heappop(heap)
        shards[s].append(i)
        heapq.heappush(heap, (load + sizes[i], s))
    for sh in shards:
        sh.sort()
    return shards


def parse_sharded(docs: Sequence[bytes], kinds: Sequence[str], parse_fn: Callable, rank: int, world: int,
                  gather: bool = True):
    """Every rank calls this with the same (docs, kinds). `parse_fn(docs, kinds) -> list` runs on
    this rank's shard (FileParser.parse_many bound to this rank's GPU). Returns the full result
    list in input order on every rank when `gather`, else only this rank's (indices, results)."""
    shards = partition([len(d) for d in docs], world)
    mine = shards[rank]
    res = parse_fn([docs[i] for i in mine], [kinds[i] for i in mine]) if mine else []
    if not gather or world == 1:
        if world == 1:
            out = [None] * len(docs)
            for i, r in zip(mine, res):
                out[i] = r
            return out
        return mine, res
    import torch.distributed as dist
    gathered = [None] * world
    dist.all_gather_object(gathered, (mine, res))
    out = [None] * len(docs)
    for idx, rs in gathered:
        for i, r in zip(idx, rs):
            out[i] = r
    return out
