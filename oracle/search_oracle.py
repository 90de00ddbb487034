"""TEST INFRASTRUCTURE ONLY - never imported by the product path.

CPU restatement of what the consumer's search returns for a collection created with Distance::Cosine
(/root/reference/services/vector_memory_service/src/main.rs:34-37 collection, :140-202 upsert, :261-286 search,
:319-403 results in descending score order). The arithmetic lives in the Qdrant server, which is not part of the
reference tree (the service only holds the client crate `qdrant-client`, vector_memory_service/Cargo.toml); restated
from its published source (lib/segment/src/spaces/simple.rs): cosine_preprocess divides a vector by its f32 length
unless its SQUARED length (is_length_zero_or_normalized) is < f32::EPSILON or within 1e-6 of 1, on insert and on the query;
the score is the dot product.
Qdrant answers exactly (full scan) below its indexing threshold and approximately (HNSW) above it; the restatement is
the exact answer. The reference holds no fixture for it: "parity unpinned" for the search row.
Scores are accumulated in float64 here, so that the checker does not depend on a summation order; ties (equal
scores) are ordered by id like the CUDA kernels order them.
"""
import numpy as np


def cosine_preprocess(v: np.ndarray) -> np.ndarray:
    v = np.asarray(v, dtype=np.float32)
    sq = np.sum(v.astype(np.float64) ** 2, axis=-1, keepdims=True).astype(np.float32)
    length = np.sqrt(sq.astype(np.float64)).astype(np.float32)
    keep = (sq < np.finfo(np.float32).eps) | (np.abs(sq - np.float32(1.0)) <= np.float32(1e-6))
    with np.errstate(divide="ignore", invalid="ignore"):
        out = (v / length).astype(np.float32)
    return np.where(keep, v, out)


def scores(vectors: np.ndarray, queries: np.ndarray) -> np.ndarray:
    """[q, n] float64 cosine scores of normalised f32 operands."""
    a = cosine_preprocess(vectors).astype(np.float64)
    b = cosine_preprocess(np.atleast_2d(queries)).astype(np.float64)
    return b @ a.T


def search(vectors: np.ndarray, queries: np.ndarray, top_k: int):
    """-> ids [q, top_k] int64 (-1 padded), scores [q, top_k] float64 (-inf padded), best first, ties by id."""
    s = scores(vectors, queries)
    q, n = s.shape
    ids = np.full((q, top_k), -1, dtype=np.int64)
    out = np.full((q, top_k), -np.inf)
    for i in range(q):
        order = np.lexsort((np.arange(n), -s[i]))[:top_k]
        ids[i, :len(order)] = order
        out[i, :len(order)] = s[i, order]
    return ids, out
