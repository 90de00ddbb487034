"""f1, the host ingest front-end (CPU only): b200_ingest_split_sentences against the restatement of main.rs:28-63 in
oracle/ingest_oracle.py, and b200_ingest_encode against the `tokenizers` library (the crate the reference links,
embedding_generator.rs:73-99, :160-187) - ids, masks and type ids bit-exact."""
import os
import random

import numpy as np
import pytest

tokenizers = pytest.importorskip("tokenizers")
from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, processors, trainers  # noqa: E402

from oracle import ingest_oracle  # noqa: E402

CORPUS = (
    "The quick brown fox jumps over the lazy dog. Sentence embeddings are computed in batches of thirty-two!\n"
    "Is tokenisation really the bottleneck? Naïve café façade déjà vu; Ångström, Zürich, Škoda, "
    "İstanbul, ǅ, straße.\n"
    "Ελληνικά κείμενα με τόνους: "
    "Αθήνα. Русский текст: Москва, "
    "ёлка.\n"
    "中文分词测试。日本語のテキスト、カタカナとひらがな。"
    "한국어 문장도 포함합니다. 豈 更 車\n"
    "Numbers 3.14159, 1,000,000 and e-mail: someone@example.com; URLs https://example.org/a/b?c=d&e=f#g\n"
    "Emoji \U0001f600 \U0001f680 and symbols © ® ™ € £ ¥ § ¶ † ‡ • … — – "
    "‘quotes’ “double” «guillemets» ¿qué? ¡hola!\n")
SPECIALS = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"]


def _finish(tok):
    tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    tok.post_processor = processors.TemplateProcessing(
        single="[CLS] $A [SEP]", pair="[CLS] $A [SEP] $B:1 [SEP]:1",
        special_tokens=[("[CLS]", tok.token_to_id("[CLS]")), ("[SEP]", tok.token_to_id("[SEP]"))])
    return tok


def _trained(lowercase=True, strip=None, chinese=True, clean=True):
    tok = Tokenizer(models.WordPiece(unk_token="[UNK]"))
    tok.normalizer = normalizers.BertNormalizer(clean_text=clean, handle_chinese_chars=chinese, strip_accents=strip, lowercase=lowercase)
    tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    tok.train_from_iterator(CORPUS.split("\n") * 3, trainers.WordPieceTrainer(vocab_size=600, special_tokens=SPECIALS, show_progress=False))
    return _finish(tok)


def _every_char(cps):
    """A vocabulary with every character as a word and as a continuation: ids tell the normalised characters apart."""
    words = list(SPECIALS)
    seen = set(words)
    for c in cps:
        for w in (chr(c), "##" + chr(c)):
            if w not in seen:
                seen.add(w)
                words.append(w)
    tok = Tokenizer(models.WordPiece({w: i for i, w in enumerate(words)}, unk_token="[UNK]"))
    tok.normalizer = normalizers.BertNormalizer(lowercase=True)
    return _finish(tok)


def _library(tok, sentences, P):
    """configure_tokenizer + tokenize_fixed of the host mirror = embedding_generator.rs:73-99, :160-187"""
    pad = tok.token_to_id("[PAD]")
    tok.enable_padding(direction="right", pad_id=0 if pad is None else pad, pad_type_id=0, pad_token="[PAD]", length=P)
    tok.enable_truncation(max_length=P, stride=0, strategy="longest_first", direction="right")
    enc = tok.encode_batch(list(sentences), add_special_tokens=True)
    return tuple(np.array([getattr(e, k) for e in enc], dtype=np.uint32) for k in ("ids", "attention_mask", "type_ids"))


def _same(pkg, tok, tmp_path, sentences, P, threads=0):
    path = os.path.join(tmp_path, "tokenizer.json")
    tok.save(path)
    ing = pkg.abi.Ingest(path)
    want = _library(tok, sentences, P)
    ids, mask, tt, lengths = ing.encode(sentences, P, threads)
    for w, g, name in zip(want, (ids, mask, tt), ("ids", "mask", "type_ids")):
        bad = np.nonzero((w != g).any(1))[0]
        assert len(bad) == 0, f"{name}: {len(bad)} rows differ, first {sentences[bad[0]]!r}: {w[bad[0]][:24]} vs {g[bad[0]][:24]}"
    assert np.array_equal(lengths, want[1].sum(1))
    return ing


_POOLS = [(0x20, 0x7E)] * 6 + [(0xA0, 0x24F), (0x300, 0x36F), (0x370, 0x3FF), (0x400, 0x4FF), (0x590, 0x6FF), (0x900, 0x97F),
                               (0x2000, 0x206F), (0x3000, 0x30FF), (0x4E00, 0x4FFF), (0xAC00, 0xAD00), (0xF900, 0xFA2F),
                               (0xFF00, 0xFFEF), (0x1F600, 0x1F64F), (0x0, 0x1F), (0x1D400, 0x1D4FF), (0xE000, 0xE010), (0x80, 0x9F)]


def _fuzz(n, seed):
    rng = random.Random(seed)
    words = CORPUS.split()
    out = []
    for _ in range(n):
        parts = []
        for _ in range(rng.randint(0, 40)):
            r = rng.random()
            if r < 0.5:
                parts.append(rng.choice(words))
            elif r < 0.6:
                parts.append(rng.choice(["[SEP]", "[CLS]", "[MASK]", "[PAD]", "[UNK]", "[SE", "##s", "#", "[", "]"]))
            else:
                lo, hi = rng.choice(_POOLS)
                parts.append("".join(chr(c) for c in (rng.randint(lo, hi) for _ in range(rng.randint(1, 12))) if not 0xD800 <= c <= 0xDFFF))
            parts.append(rng.choice([" ", "", " ", "\t", " ", "\n", "  "]))
        out.append("".join(parts))
    return out


EDGE = ["", "   ", "...", "a" * 120, "x" * 100, "x" * 101, "é" * 100, "é" * 101, "x [SEP] y [MASK]z", "tab\there\x00nul\x07bell",
        "ＡＢＣ ｆｕｌｌ", "é å", "[CLS][CLS]", "[PAD", "İstanbul ǅ ß ẞ ſ",
        "한국어", "豈更車", "á̧b", chr(0xFFFD) + chr(0xFEFF) + chr(0x200B) + chr(0xAD), "é ḉ Å Ω ̈́"]


@pytest.mark.parametrize("P", [2, 3, 8, 32, 128])
def test_encode_matches_the_library_on_corpus_and_edges(pkg, tmp_path, P):
    sentences = [s for line in CORPUS.split("\n") for s in line.split(". ")] + EDGE
    _same(pkg, _trained(), str(tmp_path), sentences, P)


def test_encode_matches_the_library_on_fuzzed_text(pkg, tmp_path):
    sentences = _fuzz(6000, 7)
    ing = _same(pkg, _trained(), str(tmp_path), sentences, 48)
    one = ing.encode(sentences, 48, threads=1)
    many = ing.encode(sentences, 48, threads=5)
    assert all(np.array_equal(a, b) for a, b in zip(one, many))  # the thread count never changes the rows


@pytest.mark.parametrize("kw", [dict(lowercase=False, strip=None), dict(lowercase=True, strip=True, chinese=False),
                                dict(lowercase=False, strip=False, clean=False)])
def test_encode_matches_the_library_for_other_normalizer_settings(pkg, tmp_path, kw):
    _same(pkg, _trained(**kw), str(tmp_path), _fuzz(3000, 11) + EDGE, 32)


def test_every_code_point_normalises_like_the_library(pkg, tmp_path):
    """'a' + c + 'b' for every code point of planes 0-2 (and the tag block) with a vocabulary that holds every character:
    clean_text, CJK spacing, NFD + accent stripping, lowercasing, Hangul decomposition, punctuation and whitespace
    classes all show up in the ids."""
    cps = [c for c in range(0x30000) if not 0xD800 <= c <= 0xDFFF] + list(range(0xE0000, 0xE0200))
    _same(pkg, _every_char(cps), str(tmp_path), ["a" + chr(c) + "b" for c in cps], 12)


def test_every_char_vocabulary_on_fuzzed_text(pkg, tmp_path):
    cps = [c for lo, hi in _POOLS for c in range(lo, hi + 1)] + list(range(0x1100, 0x1200))
    _same(pkg, _every_char(cps), str(tmp_path), _fuzz(4000, 23) + EDGE, 64)


def test_bert_processing_and_missing_pad(pkg, tmp_path):
    """BertProcessing post-processor; no [PAD] in the vocabulary -> pad id 0 (embedding_generator.rs:76-82)."""
    vocab = {w: i for i, w in enumerate(["the", "[UNK]", "[CLS]", "[SEP]", "cat", "##s", "."])}
    tok = Tokenizer(models.WordPiece(vocab, unk_token="[UNK]"))
    tok.normalizer = normalizers.BertNormalizer(lowercase=True)
    tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    tok.post_processor = processors.BertProcessing(("[SEP]", 3), ("[CLS]", 2))
    ing = _same(pkg, tok, str(tmp_path), ["The cats.", "dog", "cats " * 30], 10)
    assert ing.pad_id == 0 and ing.vocab_size == 7


def test_unsupported_tokenizers_fail_loudly(pkg, tmp_path):
    tok = Tokenizer(models.BPE())
    p = os.path.join(tmp_path, "bpe.json")
    tok.save(p)
    with pytest.raises(pkg.abi.B200EmbedError, match="WordPiece"):
        pkg.abi.Ingest(p)
    with pytest.raises(pkg.abi.B200EmbedError, match="cannot open"):
        pkg.abi.Ingest(os.path.join(tmp_path, "missing.json"))
    tok = _trained()
    tok.add_tokens(["newword"])  # a normalized, non-special added token
    p = os.path.join(tmp_path, "added.json")
    tok.save(p)
    with pytest.raises(pkg.abi.B200EmbedError, match="added token"):
        pkg.abi.Ingest(p)
    ing = _same(pkg, _trained(), str(tmp_path), ["a"], 4)
    with pytest.raises(pkg.abi.B200EmbedError, match="seq_len"):
        ing.encode(["a"], 1)


NBSP, IDSP, NEL, US, VT, FS = chr(0xA0), chr(0x3000), chr(0x85), chr(0x1F), chr(0x0B), chr(0x1C)
SPLIT_CASES = [
    "Hello world. How are you? Fine!", "No punctuation at all", "  leading and trailing   spaces .  ", "Dr. Smith went home... Really?!",
    "a.b.c", ".", "?!.", "tabs\tand\nnewlines\r\nand nbsp" + NBSP + IDSP + "ideographic separator. Nextline" + NEL + "here", "ends without stop. tail",
    "unit" + US + "separator is not whitespace. " + VT + " vertical tab is", "Unicode sentence" + chr(0x3002) + " Full-width stop is not a cut. ok",
    "3.14 is pi. v1.2.3!", " . . . ", "x" * 1000 + ".",
]


@pytest.mark.parametrize("raw", SPLIT_CASES)
def test_split_sentences_matches_the_restatement(pkg, raw):
    cleaned, begin, end = pkg.abi.split_sentences(raw)
    want_cleaned, want = ingest_oracle.split_sentences(raw)
    assert cleaned.decode() == want_cleaned
    assert [cleaned[b:e].decode() for b, e in zip(begin, end)] == want


def test_split_sentences_fuzz_and_empty(pkg):
    for raw in ("", "   ", "\n\t" + IDSP):
        with pytest.raises(ValueError, match="Cleaned text is empty"):  # main.rs:33-39
            pkg.abi.split_sentences(raw)
    rng = random.Random(5)
    alphabet = list("abc XYZ.?!,;\t\n  ") + [IDSP, NBSP, FS, NEL, chr(0xE9), chr(0x4E2D), chr(0x1F600), "  ", ". ", "? ", "..."]
    for _ in range(400):
        raw = "".join(rng.choice(alphabet) for _ in range(rng.randint(1, 200)))
        want_cleaned, want = ingest_oracle.split_sentences(raw)
        if not want_cleaned:
            continue
        cleaned, begin, end = pkg.abi.split_sentences(raw)
        assert cleaned.decode() == want_cleaned
        assert [cleaned[b:e].decode() for b, e in zip(begin, end)] == want


def test_split_then_encode_feeds_the_embedding_call(pkg, tmp_path):
    """The two calls chained the way process_raw_text_message chains them (main.rs:28-95): sentence ranges of the cleaned
    text go straight into the encoder; rows equal the library's on the sentence strings."""
    tok = _trained()
    path = os.path.join(tmp_path, "tokenizer.json")
    tok.save(path)
    ing = pkg.abi.Ingest(path)
    raw = CORPUS * 20
    cleaned, begin, end = pkg.abi.split_sentences(raw)
    ids, mask, tt, lengths = ing.encode_ranges(cleaned, begin, end, 32)
    want = _library(tok, ingest_oracle.split_sentences(raw)[1], 32)
    assert np.array_equal(ids, want[0]) and np.array_equal(mask, want[1]) and np.array_equal(tt, want[2])
    assert ids.dtype == np.uint32 and ids.shape == (len(begin), 32) and (lengths >= 2).all()


def test_split_sentences_of_a_long_text_is_chunked_across_threads(pkg):
    """Texts above 256 KB per core are cleaned in chunks (cut at ASCII whitespace) on several threads: same cleaned text,
    same sentences as the sequential restatement, including sentences and whitespace runs that straddle chunk borders."""
    rng = random.Random(11)
    pieces = ["word", "Sentence end.", "what?", "yes!", "...", "caf" + chr(0xE9), chr(0x4E2D) + chr(0x6587), "a.b", "x" * 50,
              NBSP, IDSP, "\t", "\n\n", "   ", " ", " ", " ", NEL, "tail"]
    raw = "".join(rng.choice(pieces) + rng.choice(["", " ", " ", "  ", "\n"]) for _ in range(400000))
    assert len(raw.encode()) > 8 * 256 * 1024
    cleaned, begin, end = pkg.abi.split_sentences(raw)
    want_cleaned, want = ingest_oracle.split_sentences(raw)
    assert cleaned.decode() == want_cleaned
    assert len(begin) == len(want)
    got = [cleaned[b:e].decode() for b, e in zip(begin, end)]
    assert got == want


def test_packed_encoding_is_the_padded_one_without_the_padding(pkg, tmp_path):
    """b200_ingest_encode_packed (the input of b200_embed_forward_varlen): the real tokens of the padded rows back to back,
    cu_seqlens = prefix sums of the lengths; a too-small output only reports the total; thread count does not matter."""
    tok = _trained()
    path = os.path.join(tmp_path, "tokenizer.json")
    tok.save(path)
    ing = pkg.abi.Ingest(path)
    sentences = _fuzz(3000, 5) + EDGE
    P = 40
    ids, mask, tt, lengths = ing.encode(sentences, P)
    enc = [s.encode("utf-8") for s in sentences]
    end = np.cumsum([len(b) for b in enc], dtype=np.int64)
    begin = np.concatenate([np.zeros(1, dtype=np.int64), end[:-1]])
    text = b"".join(enc)
    pi, pt, cu = ing.encode_ranges_packed(text, begin, end, P)
    assert cu.dtype == np.int32 and cu[0] == 0 and np.array_equal(np.diff(cu), lengths) and cu[-1] == len(pi) == len(pt)
    assert np.array_equal(pi, ids[mask == 1]) and np.array_equal(pt, tt[mask == 1])
    # a small caller buffer: the wrapper's first call only learns the total, the second fills
    small = (np.empty(10, np.uint32), np.empty(10, np.uint32))
    pi2, pt2, cu2 = ing.encode_ranges_packed(text, begin, end, P, threads=3, out=small)
    assert np.array_equal(pi2, pi) and np.array_equal(pt2, pt) and np.array_equal(cu2, cu)
    pi3, _, cu3 = ing.encode_ranges_packed(text, begin[:0], end[:0], P)
    assert len(pi3) == 0 and cu3.tolist() == [0]
