"""Synthetic tokenised input of the shapes BASELINE.json names.

There is no network (no tokenizer.json, no corpus), so throughput runs feed token ids with the
layout `tokenizers` hands the reference (embedding_generator.rs:160-187): `[CLS]` first, `[SEP]`
as the last real token, right padding with id 0 / mask 0 / type 0 up to the padded length.
"""
from __future__ import annotations

import numpy as np


def special_ids(vocab_size: int):
    """(cls, sep, first_ordinary_id): BERT's 101/102 when the vocabulary is big enough."""
    return (101, 102, 1000) if vocab_size > 2000 else (1, 2, 3)


def padded_batch(vocab_size: int, lengths, pad_to: int, seed: int = 42):
    """ids, mask, type_ids as [N, pad_to] uint32 for sentences of the given real lengths (>= 2)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    assert lengths.min() >= 2 and lengths.max() <= pad_to
    n = len(lengths)
    cls, sep, lo = special_ids(vocab_size)
    rng = np.random.default_rng(seed)
    ids = rng.integers(lo, vocab_size, size=(n, pad_to), dtype=np.uint32)
    pos = np.arange(pad_to)[None, :]
    mask = (pos < lengths[:, None]).astype(np.uint32)
    ids[:, 0] = cls
    ids[np.arange(n), lengths - 1] = sep
    ids *= mask
    return ids, mask, np.zeros_like(ids)


def pack(ids, mask, type_ids):
    """Padded [N,P] -> packed varlen (ids, type_ids, cu_seqlens int32[N+1]); mask must be a prefix."""
    lengths = mask.sum(axis=1).astype(np.int64)
    cu = np.zeros(len(lengths) + 1, dtype=np.int32)
    np.cumsum(lengths, out=cu[1:])
    sel = mask.astype(bool)
    return np.ascontiguousarray(ids[sel]), np.ascontiguousarray(type_ids[sel]), cu


def lengths_dense(n: int, s: int):
    return np.full(n, s, dtype=np.int64)


def lengths_uniform(n: int, lo: int, hi: int, seed: int = 7):
    return np.random.default_rng(seed).integers(lo, hi + 1, size=n)


def lengths_loguniform(n: int, lo: int = 8, hi: int = 512, seed: int = 7):
    """config c3: ragged lengths log-uniform on [lo, hi]."""
    u = np.random.default_rng(seed).uniform(np.log(lo), np.log(hi), size=n)
    return np.clip(np.round(np.exp(u)), lo, hi).astype(np.int64)


def lengths_corpus(n: int, seed: int = 2026):
    """config c5 'seq128' corpus: clip(round(LogNormal(ln 24, 0.6)), 4, 128)."""
    x = np.random.default_rng(seed).lognormal(np.log(24.0), 0.6, size=n)
    return np.clip(np.round(x), 4, 128).astype(np.int64)
