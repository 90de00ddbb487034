"""f4: the resident cosine index (b200_index_*) against oracle/search_oracle.py - the exact answer Qdrant's
Distance::Cosine search gives (vector_memory_service/src/main.rs:34-37, :261-286)."""
import os

import numpy as np
import pytest

from oracle import search_oracle

pytestmark = pytest.mark.gpu


def _check(ids, scores, vectors, queries, k, tol=2e-6):
    """Scores within tol of the f64 oracle at the returned ids; nothing better than the returned k-th was left out
    (beyond tol); descending order; ids valid and distinct."""
    s = search_oracle.scores(vectors, queries)
    want_ids, want_sc = search_oracle.search(vectors, queries, k)
    n = vectors.shape[0]
    for i in range(len(ids)):
        m = min(k, n)
        got = ids[i, :m]
        assert (got >= 0).all() and (got < n).all() and len(set(got.tolist())) == m
        assert (ids[i, m:] == -1).all() and np.isneginf(scores[i, m:]).all()
        assert np.abs(scores[i, :m] - s[i, got]).max() <= tol
        assert (np.diff(scores[i, :m]) <= 0).all()
        assert scores[i, m - 1] >= want_sc[i, m - 1] - tol  # the k-th best of the oracle is not better than ours
        # wherever the oracle's neighbours are separated by more than the tolerance the ids agree exactly
        gap = np.abs(np.diff(np.concatenate([want_sc[i, :m], [s[i, np.argsort(-s[i])[m]] if n > m else -np.inf]])))
        sure = np.concatenate([[True], gap[:-1] > 2 * tol]) & (gap > 2 * tol)
        assert (got[sure] == want_ids[i, :m][sure]).all()


@pytest.mark.parametrize("dim", [128, 384, 768, 1024])
@pytest.mark.parametrize("n,q,k", [(1, 1, 1), (37, 2, 5), (5000, 5, 10), (70001, 3, 100), (9000, 2, 48), (30000, 1, 1024)])
def test_search_matches_oracle(pkg, built, dim, n, q, k):
    rng = np.random.default_rng(dim + n)
    vectors = rng.standard_normal((n, dim)).astype(np.float32) * rng.uniform(0.1, 5.0, (n, 1)).astype(np.float32)
    queries = rng.standard_normal((q, dim)).astype(np.float32)
    with pkg.abi.Index(dim) as ix:
        ix.add(vectors[: n // 2])
        ix.add(vectors[n // 2:])  # growth + append keep the ids
        assert len(ix) == n
        ids, scores = ix.search(queries, k)
    _check(ids, scores, vectors, queries, k)


def test_exact_ids_and_tie_order_when_the_arithmetic_is_exact(pkg, built):
    """A corpus of signed one-hot rows (length exactly 1, so cosine_preprocess leaves them alone) and unit-length queries
    with a few components that are multiples of 1/8: every product and every sum is exact in f32 whatever the order,
    and duplicates abound. The ids must then match the oracle bit for bit, equal scores ordered by id."""
    dim, n, k = 256, 20000, 64
    rng = np.random.default_rng(1)
    hot = rng.integers(0, dim, n)
    sign = rng.choice([-1.0, 1.0], n).astype(np.float32)
    vectors = np.zeros((n, dim), dtype=np.float32)
    vectors[np.arange(n), hot] = sign
    q = np.zeros((2, dim), dtype=np.float32)
    pos = rng.choice(dim, 11, replace=False)
    q[0, pos[:4]] = 3 / 8   # 4 * 9/64 + 7 * 4/64 = 1
    q[0, pos[4:]] = 2 / 8
    q[1, 3] = 1.0
    with pkg.abi.Index(dim, capacity_hint=n) as ix:
        ix.add(vectors)
        ids, scores = ix.search(q, k)
    for i in range(2):
        want = sign * q[i, hot]
        order = np.lexsort((np.arange(n), -want.astype(np.float64)))[:k]
        assert np.array_equal(ids[i], order)
        assert np.array_equal(scores[i], want[order])


def test_small_index_padding_and_errors(pkg, built):
    with pkg.abi.Index(384) as ix:
        ids, scores = ix.search(np.ones((2, 384), np.float32), 4)
        assert (ids == -1).all() and np.isneginf(scores).all()
        v = np.eye(3, 384, dtype=np.float32)
        ix.add(v)
        ids, scores = ix.search(v[1], 5)
        assert ids[0].tolist() == [1, 0, 2, -1, -1] and scores[0, 0] == 1.0 and scores[0, 1] == 0.0
        with pytest.raises(pkg.abi.B200EmbedError, match="top_k"):
            ix.search(v[1], 0)
        with pytest.raises(pkg.abi.B200EmbedError, match="top_k"):
            ix.search(v[1], 2000)
        with pytest.raises(ValueError):
            ix.add(np.zeros((2, 100), np.float32))
    with pytest.raises(pkg.abi.B200EmbedError, match="dimension"):
        pkg.abi.Index(100)


def test_zero_and_unit_vectors_are_left_alone(pkg, built):
    """cosine_preprocess: a zero vector stays zero (score 0), a unit vector is not rescaled."""
    dim = 128
    v = np.zeros((4, dim), dtype=np.float32)
    v[1, 0] = 1.0
    v[2, :2] = [3.0, 4.0]
    v[3, 5] = 1e-9  # shorter than f32::EPSILON: unchanged, scores stay tiny
    q = np.zeros(dim, dtype=np.float32)
    q[0] = 2.0
    q[5] = 2.0
    with pkg.abi.Index(dim) as ix:
        ix.add(v)
        ids, scores = ix.search(q, 4)
    _check(ids, scores, v, q[None, :], 4)


def test_embeddings_from_the_path_feed_the_index(pkg, built):
    """Embed sentences on the GPU, add the device buffer straight to the index, search with one of them: the sentence
    finds itself first with score 1."""
    import torch
    shape = pkg.weights.TINY
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [5, 9, 17, 33, 12, 7, 64, 21], 64, seed=3)
    with pkg.Embedder.random_init(shape, seed=7) as emb, pkg.abi.Index(shape.hidden_size) as ix:
        out = emb.forward(ids, mask, tt)
        d_out = torch.from_numpy(out).cuda()
        ix.add_device(d_out.data_ptr(), out.shape[0])
        got_ids, got_sc = ix.search(out[4], 3)
    assert got_ids[0, 0] == 4 and abs(got_sc[0, 0] - 1.0) <= 2e-6
    _check(got_ids, got_sc, out, out[4:5], 3)


def test_selection_with_many_survivors_falls_back_to_the_full_sort(pkg, built):
    """The k <= 256 selection keeps what is at least as good as the k-th best of the 256 per-thread bests of a 4096-block.
    When the good candidates all sit in the slots of k threads (slot i belongs to thread i % 256) there are 16 k
    survivors, more than the 1024 the filter sorts, and the block is sorted as a whole instead: same answer."""
    dim, n, k = 128, 4096 * 3 + 500, 100
    rng = np.random.default_rng(9)
    theta = np.full(n, 2.0)  # cos(2.0) < 0: never in the answer
    good = np.nonzero((np.arange(n) % 256) < k)[0]
    theta[good] = 0.01 + 0.001 * rng.permutation(len(good))
    vectors = np.zeros((n, dim), dtype=np.float32)
    vectors[:, 0] = np.cos(theta)
    vectors[:, 1] = np.sin(theta)
    q = np.zeros((1, dim), dtype=np.float32)
    q[0, 0] = 1.0
    with pkg.abi.Index(dim) as ix:
        ix.add(vectors)
        ids, scores = ix.search(q, k)
    want_ids, want_sc = search_oracle.search(vectors, q, k)
    assert np.array_equal(ids, want_ids)
    assert np.abs(scores - want_sc).max() <= 2e-6


def test_save_and_load_give_the_same_answers(pkg, built, tmp_path):
    rng = np.random.default_rng(4)
    vectors = rng.standard_normal((30000, 256)).astype(np.float32)
    queries = rng.standard_normal((3, 256)).astype(np.float32)
    path = str(tmp_path / "corpus.b200idx")
    with pkg.abi.Index(256) as ix:
        ix.add(vectors)
        want = ix.search(queries, 20)
        ix.save(path)
    assert os.path.getsize(path) == 8 + 4 + 8 + vectors.size * 4
    with pkg.abi.Index.load(path) as ix2:
        assert len(ix2) == 30000 and ix2.dim == 256
        got = ix2.search(queries, 20)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])  # bit for bit
        ix2.add(vectors[:5])  # a loaded index keeps growing
        assert len(ix2) == 30005
    with open(path, "r+b") as f:
        f.truncate(1000)
    with pytest.raises(pkg.abi.B200EmbedError, match="truncated"):
        pkg.abi.Index.load(path)
    with pytest.raises(pkg.abi.B200EmbedError, match="cannot open"):
        pkg.abi.Index.load(str(tmp_path / "missing"))


@pytest.mark.parametrize("dim,n,k", [(384, 120000, 10), (768, 50000, 64), (128, 9000, 256), (256, 300, 5), (1024, 20000, 1)])
def test_two_stage_search_gives_the_same_bits(pkg, built, dim, n, k):
    """With the bf16 copy enabled the scan reads half the bytes and the candidates are rescored from the f32 rows: ids and
    scores must be the very same bits as the one-stage search, for rows added before and after enabling it."""
    rng = np.random.default_rng(dim + k)
    vectors = rng.standard_normal((n, dim)).astype(np.float32)
    queries = rng.standard_normal((5, dim)).astype(np.float32)
    with pkg.abi.Index(dim) as plain, pkg.abi.Index(dim) as fast:
        plain.add(vectors)
        fast.add(vectors[: n // 3])
        fast.enable_prescan()
        fast.add(vectors[n // 3:])  # grows the matrix and its copy
        want = plain.search(queries, k)
        got = fast.search(queries, k)  # 5 queries: a pass of four (one stage) and a pass of one (two stages if k <= 64)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        for i in range(5):  # one query per call, like the reference's search requests
            one = fast.search(queries[i], k)
            assert np.array_equal(one[0][0], want[0][i]) and np.array_equal(one[1][0], want[1][i])
        assert fast.exact_fallbacks == 0
    _check(got[0], got[1], vectors, queries, k)


def test_two_stage_search_falls_back_when_scores_crowd(pkg, built):
    """Thousands of rows within 1e-3 of each other: the bf16 stage cannot separate them, the bound says so, and the search
    is repeated on the f32 matrix - still the exact answer."""
    dim, n, k = 256, 20000, 10
    rng = np.random.default_rng(3)
    base = rng.standard_normal(dim).astype(np.float32)
    vectors = (base[None, :] + 1e-4 * rng.standard_normal((n, dim))).astype(np.float32)
    q = base[None, :].copy()
    with pkg.abi.Index(dim) as plain, pkg.abi.Index(dim) as fast:
        plain.add(vectors)
        fast.add(vectors)
        fast.enable_prescan()
        want = plain.search(q, k)
        got = fast.search(q, k)
        assert fast.exact_fallbacks == 1
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
