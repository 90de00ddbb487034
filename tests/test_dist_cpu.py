"""CPU, world_size 2 over gloo: the N > 1 host logic -- shard plan, per-rank work, gather to rank 0, order
restoration, max-over-ranks timing. The per-rank "embedding" is a stand-in computed with numpy (the product path
needs a GPU); what is under test is the plumbing around it."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_embed(ids_row, hidden):
    v = np.cos(np.arange(hidden)[None, :] * 0.01 * (ids_row.sum() % 97 + 1)) * (1 + (ids_row != 0).sum())
    return v.astype(np.float32)[0]


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from __graft_entry__ import load_package
    pkg = load_package()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shape = pkg.weights.TINY
    lengths = pkg.synth.lengths_corpus(500, seed=2026).clip(4, 64)
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 64, seed=42)
    plan = pkg.sharding.plan_shards(lengths, world, token_budget=700, window=128, shape=shape)
    parts = [(b, np.stack([_fake_embed(ids[i], shape.hidden_size) for i in b])) for b in plan[rank]]
    t_local = torch.tensor([0.010 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t_local, op=dist.ReduceOp.MAX)  # bench.py: time = max over ranks
    gathered = [None] * world
    dist.all_gather_object(gathered, parts)
    if rank == 0:
        flat = [p for rank_parts in gathered for p in rank_parts]
        out = pkg.sharding.scatter_back(len(lengths), shape.hidden_size, flat)
        want = np.stack([_fake_embed(ids[i], shape.hidden_size) for i in range(len(lengths))])
        q.put((bool(np.array_equal(out, want)), float(t_local.item()), pkg.sharding.imbalance(plan, lengths, shape),
               sorted(len(r) for r in plan)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shard_gather_restores_order():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    ok, tmax, imb, sizes = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert ok and abs(tmax - 0.020) < 1e-12
    assert imb < 1.15 and min(sizes) > 0


def test_plan_covers_every_sentence_once_and_respects_budget(pkg):
    lengths = pkg.synth.lengths_loguniform(3000, 8, 512, seed=7)
    plan = pkg.sharding.plan_shards(lengths, 8, token_budget=4096, shape=pkg.weights.MINILM_L6)
    allidx = np.concatenate([b for r in plan for b in r])
    assert sorted(allidx.tolist()) == list(range(3000))
    assert all(lengths[b].sum() <= 4096 or len(b) == 1 for r in plan for b in r)
    assert pkg.sharding.imbalance(plan, lengths, pkg.weights.MINILM_L6) < 1.05
    with pytest.raises(ValueError):
        pkg.sharding.scatter_back(3, 4, [(np.array([0, 1]), np.zeros((2, 4), np.float32))])


def _search_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from __graft_entry__ import load_package
    from oracle import search_oracle
    pkg = load_package()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)
    vectors = rng.standard_normal((1001, 64)).astype(np.float32)
    vectors[700:720] = vectors[10:30]  # duplicates across the shards: ties must come out ordered by global id
    queries = rng.standard_normal((3, 64)).astype(np.float32)
    queries[2] = vectors[15]
    bounds = [0, 333, 1001][: world + 1] if world == 2 else np.linspace(0, 1001, world + 1).astype(int).tolist()
    lo, hi = bounds[rank], bounds[rank + 1]

    def local_search(qs, k):  # stand-in for this rank's b200_index (needs a GPU): the oracle on the rank's rows
        ids, sc = search_oracle.search(vectors[lo:hi], qs, k)
        return ids, sc.astype(np.float32)

    ids, sc = pkg.sharding.search_sharded(local_search, lo, queries, 40)
    want_ids, want_sc = search_oracle.search(vectors, queries, 40)
    q.put((rank, np.array_equal(ids, want_ids), float(np.abs(sc - want_sc).max()), int(ids[2, 0]), int(ids[2, 1])))
    dist.destroy_process_group()


def test_two_rank_sharded_search_merges_to_the_global_answer():
    """f4 across GPUs: rows sharded over 2 ranks, each searches its shard, one all_gather of the k-lists, merge. The
    result equals the single-index answer on every rank, duplicates across shards ordered by global id."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_search_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(30)
    for rank, same, err, first, second in res:
        assert same and err <= 1e-6
        assert (first, second) == (15, 705)  # the query's own row, then its duplicate in the other shard


def test_merge_topk_padding_and_order(pkg):
    ids_a = np.array([[4, 2, -1]]); sc_a = np.array([[0.9, 0.5, -np.inf]], np.float32)
    ids_b = np.array([[7, 9, 8]]); sc_b = np.array([[0.9, 0.5, 0.1]], np.float32)
    ids, sc = pkg.sharding.merge_topk([ids_a, ids_b], [sc_a, sc_b], 6)
    assert ids[0].tolist() == [4, 7, 2, 9, 8, -1] and np.isneginf(sc[0, 5]) and sc[0, :5].tolist() == pytest.approx([0.9, 0.9, 0.5, 0.5, 0.1])


def test_native_corpus_planner_properties(pkg, built):
    """b200_corpus_plan (the planner of the one-process multi-GPU driver): a permutation, ascending token counts, stable for
    equal lengths, every micro-batch inside both budgets, nothing lost at the budget edges."""
    rng = np.random.default_rng(0)
    for n, max_s, max_t in [(0, 4, 100), (1, 4, 100), (1000, 64, 900), (70000, 16384, 512 * 1024), (5000, 1 << 30, 128)]:
        lengths = rng.integers(1, 129, size=n).astype(np.int32)
        order, begins = pkg.abi.corpus_plan(lengths, 128, max_s, max_t)
        assert sorted(order.tolist()) == list(range(n))
        ls = lengths[order]
        assert np.all(np.diff(ls) >= 0)
        for l in np.unique(ls):
            idx = order[ls == l]
            assert np.all(np.diff(idx) > 0)  # stable
        assert begins[0] == 0 and begins[-1] == n and (n == 0 or np.all(np.diff(begins) > 0))
        for a, b in zip(begins[:-1], begins[1:]):
            assert b - a <= max_s and ls[a:b].sum() <= max(max_t, 128)
        # greedy cuts: a batch could not have taken the next sentence as well
        for a, b in zip(begins[:-2], begins[1:-1]):
            assert b - a == max_s or ls[a:b].sum() + ls[b] > max_t


def _greedy_tiles(lengths, seg=128, cap=128):
    """tiles the engine's packing kernel makes of sentences sent in this order (csrc/elementwise.cu pack_tiles_kernel)"""
    n = 0
    for s0 in range(0, len(lengths), seg):
        base = cnt = 0
        for l in lengths[s0:s0 + seg]:
            if base + l > cap and cnt > 0:
                n, base, cnt = n + 1, 0, 0
            base, cnt = base + l, cnt + 1
        n += 1
    return n


def test_native_corpus_planner_packs_tiles(pkg, built):
    """b200_corpus_plan_packed (what b200_corpus_submit plans): a permutation inside the budgets; sentences of at most 128
    tokens stay in input-order micro-batches (as sets) and come out tile by tile - the packing kernel's greedy tiles are
    nearly full, far fuller than from the length order; longer sentences keep the ascending order of b200_corpus_plan."""
    rng = np.random.default_rng(3)
    lengths = np.clip(np.rint(rng.lognormal(np.log(24), 0.6, size=70000)), 4, 128).astype(np.int32)
    max_s, max_t = 16384, 512 * 1024
    order, begins = pkg.abi.corpus_plan_packed(lengths, 512, max_s, max_t)
    assert sorted(order.tolist()) == list(range(len(lengths)))
    assert begins[0] == 0 and begins[-1] == len(lengths) and np.all(np.diff(begins) > 0)
    tiles = 0
    for k, (a, b) in enumerate(zip(begins[:-1], begins[1:])):
        ls = lengths[order[a:b]]
        assert b - a <= max_s and ls.sum() <= max_t
        # input-order cut: micro-batch k is the k-th run of consecutive sentences
        assert sorted(order[a:b].tolist()) == list(range(a, b))
        tiles += _greedy_tiles(ls.tolist())
    ideal = lengths.sum() / 128
    by_length, _ = pkg.abi.corpus_plan(lengths, 512, max_s, max_t)
    assert tiles <= 1.06 * ideal, (tiles, ideal)  # 128-sentence segments cost a few per cent
    assert tiles <= 0.87 * _greedy_tiles(lengths[by_length].tolist())
    # mixed corpus: the long sentences follow, ascending, in micro-batches of their own
    lengths2 = np.concatenate([lengths[:5000], rng.integers(129, 513, size=800).astype(np.int32)])
    rng.shuffle(lengths2)
    order2, begins2 = pkg.abi.corpus_plan_packed(lengths2, 512, max_s, max_t)
    assert sorted(order2.tolist()) == list(range(len(lengths2)))
    ls2 = lengths2[order2]
    n_short = int((lengths2 <= 128).sum())
    assert np.all(ls2[:n_short] <= 128) and np.all(np.diff(ls2[n_short:]) >= 0) and n_short in begins2.tolist()
    for a, b in zip(begins2[:-1], begins2[1:]):
        assert b - a <= max_s and ls2[a:b].sum() <= max_t
