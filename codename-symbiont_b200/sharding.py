"""Host-side sharding of a sentence stream across the GPUs of one box.

The path has no exchange step: sentences are independent, each rank owns a full replica. What the host has to do
is (1) cut the stream into micro-batches that fit a token budget and cost about the same, (2) deal them to ranks so
that all ranks finish together, (3) put the results back in the order the caller gave
(`generate_sentence_embeddings` must preserve order, embedding_generator.rs:215; main.rs:108-115 zips by index).
Cost model: F(s) = L (24 H^2 s + 4 H s^2) FLOPs per sentence of s real tokens.
"""
from __future__ import annotations

import numpy as np


def plan_shards(lengths, world: int, token_budget: int, window: int = 65536, shape=None):
    """-> list over ranks of lists of index arrays (micro-batches of original sentence indices).

    Sentences are sorted by length inside windows of `window` sentences (short with short: fewer wasted
    128-row tile rows, similar cost per micro-batch), cut at `token_budget` real tokens, then dealt greedily to
    the rank with the least accumulated cost (longest-processing-time rule)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    n = len(lengths)
    if shape is not None:
        cost = np.array([shape.flops_per_sentence(int(s)) for s in np.unique(lengths)])
        cost = dict(zip(np.unique(lengths).tolist(), cost.tolist()))
        cost_of = np.vectorize(cost.get, otypes=[np.float64])
    else:
        cost_of = lambda s: s.astype(np.float64)
    batches = []
    for lo in range(0, n, window):
        idx = np.arange(lo, min(n, lo + window))
        idx = idx[np.argsort(lengths[idx], kind="stable")]
        start, tok = 0, 0
        for k, i in enumerate(idx):
            if tok + lengths[i] > token_budget and k > start:
                batches.append(idx[start:k])
                start, tok = k, 0
            tok += lengths[i]
        if start < len(idx):
            batches.append(idx[start:])
    costs = [float(cost_of(lengths[b]).sum()) for b in batches]
    load = np.zeros(world)
    out = [[] for _ in range(world)]
    for k in np.argsort(costs)[::-1]:
        r = int(np.argmin(load))
        out[r].append(batches[k])
        load[r] += costs[k]
    return out


def scatter_back(n: int, hidden: int, parts):
    """parts: iterable of (index array, [len(index), hidden] float32 array) -> [n, hidden] in the caller's order."""
    out = np.empty((n, hidden), dtype=np.float32)
    seen = np.zeros(n, dtype=bool)
    for idx, emb in parts:
        if seen[idx].any():
            raise ValueError("a sentence was embedded twice")
        out[idx] = emb
        seen[idx] = True
    if not seen.all():
        raise ValueError(f"{int((~seen).sum())} sentences were never embedded")
    return out


def imbalance(plan, lengths, shape=None) -> float:
    """max rank cost / mean rank cost (1.0 = perfect)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    f = (lambda s: shape.flops_per_sentence(int(s))) if shape is not None else float
    loads = [sum(sum(f(s) for s in lengths[b]) for b in rank) for rank in plan]
    return max(loads) / (sum(loads) / len(loads))


# ---- f4 across GPUs: every rank holds a slice of the vectors in its own b200_index --------------------------------
def merge_topk(ids_parts, scores_parts, k: int):
    """Merges per-shard results of the same queries: ids_parts / scores_parts are lists of [q, k_r] arrays holding GLOBAL
    ids (-1 = padding) and scores, best first. -> ids [q, k], scores [q, k], best first, equal scores by id, padded
    with (-1, -inf): the list a single index over all the vectors returns."""
    ids = np.concatenate([np.asarray(a, dtype=np.int64) for a in ids_parts], axis=1)
    scores = np.concatenate([np.asarray(a, dtype=np.float32) for a in scores_parts], axis=1)
    q = ids.shape[0]
    out_ids = np.full((q, k), -1, dtype=np.int64)
    out_sc = np.full((q, k), -np.inf, dtype=np.float32)
    for i in range(q):
        real = np.nonzero(ids[i] >= 0)[0]
        order = real[np.lexsort((ids[i, real], -scores[i, real].astype(np.float64)))][:k]
        out_ids[i, :len(order)] = ids[i, order]
        out_sc[i, :len(order)] = scores[i, order]
    return out_ids, out_sc


def search_sharded(local_search, id_offset: int, queries, k: int, group=None):
    """One search over an index sharded by rows across the ranks of `group` (torch.distributed). `local_search(queries,
    k) -> (ids, scores)` is this rank's b200_index search (local row ids); `id_offset` the global id of its row 0. The
    only exchange is an all_gather of the [q, k] lists (8 + 4 bytes per entry): NCCL between GPUs, gloo in the CPU tests.
    Every rank returns the merged result."""
    import torch
    import torch.distributed as dist

    ids, scores = local_search(queries, k)
    ids = np.where(ids >= 0, ids + id_offset, -1).astype(np.int64)
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t_ids, t_sc = torch.from_numpy(ids).to(dev), torch.from_numpy(np.ascontiguousarray(scores, dtype=np.float32)).to(dev)
    g_ids = [torch.empty_like(t_ids) for _ in range(world)]
    g_sc = [torch.empty_like(t_sc) for _ in range(world)]
    dist.all_gather(g_ids, t_ids, group=group)
    dist.all_gather(g_sc, t_sc, group=group)
    return merge_topk([t.cpu().numpy() for t in g_ids], [t.cpu().numpy() for t in g_sc], k)
