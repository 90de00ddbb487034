"""Host-side sharding of a sentence stream across the GPUs of one box.

The path has no exchange step: sentences are independent, each rank owns a full replica. What the host has to do
is (1) cut the stream into micro-batches that fit a token budget and cost about the same, (2) deal them to ranks so
that all ranks finish together, (3) put the results back in the order the caller gave
(`generate_sentence_embeddings` must preserve order, embedding_generator.rs:215; main.rs:108-115 zips by index).
Cost model: F(s) = L (24 H^2 s + 4 H s^2) FLOPs per sentence of s real tokens.
"""
from __future__ import annotations

import numpy as np


def plan_shards(lengths, world: int, token_budget: int, window: int = 65536, shape=None, max_batch_sentences: int = 1 << 30):
    """-> list over ranks of lists of index arrays (micro-batches of original sentence indices).

    A STATIC plan, for callers that run one process per GPU (torchrun) and cannot share a queue: sentences are sorted by
    length inside windows of `window` sentences and cut at `token_budget` real tokens (the native planner,
    b200_corpus_plan: short with short, so fewer wasted 128-row tile rows and similar cost per micro-batch), then dealt
    greedily to the rank with the least accumulated cost (longest-processing-time rule). The one-process driver
    (b200_corpus_*) needs no dealing: its GPU workers pull micro-batches from one queue. Vectorised: 10 M sentences plan in
    about a second."""
    from .abi import corpus_plan
    lengths = np.asarray(lengths, dtype=np.int64)
    n = len(lengths)
    if shape is not None:  # F(s) per distinct length, looked up through a table
        table = np.zeros(int(lengths.max()) + 1 if n else 1)
        for s in np.unique(lengths):
            table[int(s)] = shape.flops_per_sentence(int(s))
        sent_cost = table[lengths]
    else:
        sent_cost = lengths.astype(np.float64)
    batches, costs = [], []
    max_len = int(lengths.max()) if n else 1
    for lo in range(0, n, window):
        hi = min(n, lo + window)
        order, begins = corpus_plan(lengths[lo:hi], max(max_len, 1), max_batch_sentences, max(token_budget, max_len))
        csum = np.concatenate([[0.0], np.cumsum(sent_cost[lo:hi][order])])
        for k in range(len(begins) - 1):
            batches.append(order[begins[k]:begins[k + 1]].astype(np.int64) + lo)
            costs.append(csum[begins[k + 1]] - csum[begins[k]])
    load = np.zeros(world)
    out = [[] for _ in range(world)]
    for k in np.argsort(costs, kind="stable")[::-1]:
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
    if shape is not None:
        table = np.zeros(int(lengths.max()) + 1)
        for s in np.unique(lengths):
            table[int(s)] = shape.flops_per_sentence(int(s))
        cost = table[lengths]
    else:
        cost = lengths.astype(np.float64)
    loads = [sum(float(cost[b].sum()) for b in rank) for rank in plan]
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
