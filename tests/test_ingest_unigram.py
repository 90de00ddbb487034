"""f1 for the tokenizer the service actually runs (CPU only): the Unigram / SentencePiece pipeline of an XLM-R tokenizer.json
(paraphrase-multilingual-mpnet-base-v2, /root/reference/services/preprocessing_service/src/main.rs:305; set-up
embedding_generator.rs:73-99, encode_batch :160-187) in b200_ingest_*, bit-exact against the `tokenizers` library - the Python
binding of the crate the reference links. The fixture tests/golden/xlmr_like_tokenizer.json (scripts/make_unigram_fixture.py)
carries SentencePiece's real nmt_nfkc charsmap; the vocabulary is trained offline on synthetic multi-script text."""
import json
import os
import random
import time

import numpy as np
import pytest

tokenizers = pytest.importorskip("tokenizers")
from tokenizers import Tokenizer  # noqa: E402

from conftest import GOLDEN  # noqa: E402
from unigram_corpus import CORPUS, FUZZ_EXTRAS  # noqa: E402

FIXTURE = os.path.join(GOLDEN, "xlmr_like_tokenizer.json")


def _library(tok, sentences, P):
    """embedding_generator.rs:73-99: "[PAD]" is not in an XLM-R vocabulary, so the reference pads with id 0"""
    pad = tok.token_to_id("[PAD]")
    tok.enable_padding(direction="right", pad_id=0 if pad is None else pad, pad_type_id=0, pad_token="[PAD]", length=P)
    tok.enable_truncation(max_length=P, stride=0, strategy="longest_first", direction="right")
    enc = tok.encode_batch(list(sentences), add_special_tokens=True)
    return tuple(np.array([getattr(e, k) for e in enc], dtype=np.uint32) for k in ("ids", "attention_mask", "type_ids"))


@pytest.fixture(scope="module")
def pair(pkg):
    return Tokenizer.from_file(FIXTURE), pkg.abi.Ingest(FIXTURE)


def _same(pair, sentences, P, threads=0):
    tok, ing = pair
    want = _library(tok, sentences, P)
    ids, mask, tt, lengths = ing.encode(sentences, P, threads)
    for w, g, name in zip(want, (ids, mask, tt), ("ids", "mask", "type_ids")):
        bad = np.nonzero((w != g).any(1))[0]
        assert len(bad) == 0, f"{name}: {len(bad)} rows differ, first {sentences[bad[0]]!r}: {w[bad[0]][:32]} vs {g[bad[0]][:32]}"
    assert np.array_equal(lengths, want[1].sum(1))


@pytest.mark.parametrize("P", [2, 3, 16, 128, 514])
def test_unigram_matches_the_library_on_corpus_and_edges(pair, P):
    _same(pair, CORPUS, P)
    assert pair[1].pad_id == 0 and pair[1].vocab_size == 6001


def _fuzz_sentence(rng):
    pools = [(0x20, 0x7e), (0xa0, 0x17f), (0x300, 0x36f), (0x370, 0x3ff), (0x400, 0x4ff), (0x590, 0x5ff), (0x600, 0x6ff), (0x900, 0x97f),
             (0x980, 0x9ff), (0xe00, 0xe7f), (0x1100, 0x11ff), (0x2000, 0x206f), (0x20d0, 0x20ff), (0x2100, 0x218f), (0x2460, 0x24ff),
             (0x2580, 0x2582), (0x3000, 0x30ff), (0x3200, 0x33ff), (0x4e00, 0x4e40), (0xac00, 0xac40), (0xfb00, 0xfb4f), (0xfe00, 0xfe0f),
             (0xff00, 0xffef), (0x1f1e6, 0x1f1ff), (0x1f300, 0x1f340), (0x1f3fb, 0x1f3ff), (0xe0020, 0xe007f)]
    out = []
    for _ in range(rng.randint(0, 40)):
        lo, hi = rng.choice(pools)
        if rng.random() < 0.4:
            lo, hi = 0x61, 0x7a
        c = rng.randint(lo, hi)
        if 0xd800 <= c <= 0xdfff:
            continue
        out.append(chr(c))
        if rng.random() < 0.15:
            out.append(rng.choice(FUZZ_EXTRAS))
    return "".join(out)


def test_unigram_matches_the_library_on_fuzzed_text(pair):
    rng = random.Random(7)
    _same(pair, [_fuzz_sentence(rng) for _ in range(20000)], 48)


def test_every_code_point_normalises_and_segments_like_the_library(pair):
    """Every code point of planes 0-2 and 14 after a base letter, between letters, and before one: the Precompiled normalizer's
    grapheme walk (combining marks, prepended marks, controls, joiners) and its charsmap look-ups, the whitespace split, the
    Metaspace split and the unknown-piece fusing all show in the ids."""
    cps = [c for c in list(range(0, 0x30000)) + list(range(0xE0000, 0xE0200)) if not 0xD800 <= c <= 0xDFFF]
    e_acute = chr(0xE9)
    for lo in range(0, len(cps), 40000):
        part = cps[lo:lo + 40000]
        _same(pair, ["a" + chr(c) + "b" for c in part], 12)
        _same(pair, [chr(c) + e_acute for c in part], 12)
        _same(pair, [e_acute + chr(c) + chr(c) for c in part], 12)


def test_unigram_packed_and_threads(pair):
    tok, ing = pair
    sents = (CORPUS * 40)[:600]
    _same(pair, sents, 64, threads=4)
    enc = [s.encode() for s in sents]
    text = b"".join(enc)
    end = np.cumsum([len(b) for b in enc]).astype(np.int64)
    begin = np.concatenate([[0], end[:-1]]).astype(np.int64)
    ids, tt, cu = ing.encode_ranges_packed(text, begin, end, 64, threads=3)
    want = _library(tok, sents, 64)
    for i in range(len(sents)):
        n = int(want[1][i].sum())
        assert cu[i + 1] - cu[i] == n and np.array_equal(ids[cu[i]:cu[i + 1]], want[0][i][:n])


def test_unigram_is_faster_than_the_library(pair):
    tok, ing = pair
    rng = random.Random(1)
    words = [w for s in CORPUS[:6] for w in s.split()]
    sents = [" ".join(rng.choice(words) for _ in range(rng.randint(5, 30))) for _ in range(20000)]
    _library(tok, sents[:100], 128)
    t0 = time.perf_counter()
    _library(tok, sents, 128)
    t_lib = time.perf_counter() - t0
    ing.encode(sents[:100], 128)
    t0 = time.perf_counter()
    ing.encode(sents, 128)
    t_nat = time.perf_counter() - t0
    print(f"\n[unigram] library {len(sents) / t_lib:,.0f} sentences/s, native {len(sents) / t_nat:,.0f} sentences/s ({t_lib / t_nat:.1f}x)")
    assert t_nat < t_lib


def test_other_unigram_shapes_are_refused_by_name(pkg, tmp_path):
    d = json.load(open(FIXTURE))
    for key, val, msg in [("normalizer", {"type": "NFKC"}, "normalizer NFKC"),
                          ("pre_tokenizer", {"type": "Whitespace"}, "pre_tokenizer Whitespace"),
                          ("pre_tokenizer", None, "Metaspace")]:
        e = dict(d)
        e[key] = val
        p = os.path.join(tmp_path, "bad.json")
        json.dump(e, open(p, "w"))
        with pytest.raises(pkg.abi.B200EmbedError, match=msg):
            pkg.abi.Ingest(p)
    e = json.loads(json.dumps(d))
    e["model"]["byte_fallback"] = True
    p = os.path.join(tmp_path, "bf.json")
    json.dump(e, open(p, "w"))
    with pytest.raises(pkg.abi.B200EmbedError, match="byte_fallback"):
        pkg.abi.Ingest(p)
