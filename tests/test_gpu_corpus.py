"""GPU: the corpus driver (b200_corpus_*, BASELINE config 5's data path): shards of text in, embeddings out in input order
whatever micro-batch and GPU computed them, several shards in flight, every visible GPU used, errors reported per shard."""
import numpy as np
import pytest

from conftest import cos_rows, CROSS_SCHEDULE_ABS, GATE_COS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tiny_tokenizer(tmp_path_factory):
    from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, processors
    words = [f"w{i}" for i in range(400)]
    vocab = {w: i for i, w in enumerate(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "."] + words)}
    tok = Tokenizer(models.WordPiece(vocab, unk_token="[UNK]"))
    tok.normalizer = normalizers.BertNormalizer(lowercase=True)
    tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    tok.post_processor = processors.TemplateProcessing(single="[CLS] $A [SEP]", special_tokens=[("[CLS]", 2), ("[SEP]", 3)])
    p = str(tmp_path_factory.mktemp("tok") / "tokenizer.json")
    tok.save(p)
    return p, words


def make_text(words, lengths, seed):
    rng = np.random.default_rng(seed)
    return " ".join(" ".join(words[i] for i in rng.integers(0, len(words), size=int(n))) + "." for n in lengths)


@pytest.mark.parametrize("operands", ["f16", "bf16"])
def test_corpus_matches_direct_calls_in_order(pkg, built, tiny_tokenizer, operands):
    import torch
    path, words = tiny_tokenizer
    shape = pkg.weights.TINY  # vocabulary 512 >= the tokenizer's 405 ids, 64 positions
    ingest = pkg.abi.Ingest(path)
    devices = list(range(torch.cuda.device_count()))
    shards = []
    for s in range(5):  # sentence counts that are not multiples of the 2048-sentence tokeniser chunk or the batch budget
        lengths = np.random.default_rng(s).integers(1, 60, size=[5000, 3, 2048, 4097, 1][s])
        cleaned, begin, end = pkg.abi.split_sentences(make_text(words, lengths, 100 + s))
        assert len(begin) == len(lengths)
        shards.append((cleaned, begin, end))
    with pkg.abi.Corpus(shape, ingest, devices, seed=77, tokenizer_threads=3, max_batch_sentences=700, max_batch_tokens=9000,
                        operands=operands) as corpus, pkg.Embedder.random_init(shape, seed=77, operands=operands) as emb:
        outs = [np.full((len(b), shape.hidden_size), np.nan, np.float32) for _, b, _ in shards]
        tickets = [corpus.submit(c, b, e, o) for (c, b, e), o in zip(shards, outs)]  # all in flight at once
        for t in reversed(tickets):  # waited for in another order than submitted
            corpus.wait(t)
        st = corpus.stats()
        assert st["sentences"] == sum(len(b) for _, b, _ in shards) and st["gpus"] == len(devices)
        assert sum(st["per_gpu_sentences"]) == st["sentences"] and st["micro_batches"] >= 5 + 5000 // 700
        for (cleaned, begin, end), got in zip(shards, outs):
            ids, tt, cu = ingest.encode_ranges_packed(cleaned, begin, end, shape.max_position_embeddings)
            want = emb.forward_varlen(ids, tt, cu, shape.max_position_embeddings)
            assert np.isfinite(got).all()
            # the same sentences in other batch compositions: two schedules of the same 16-bit math, row for row
            assert np.abs(got - want).max() <= CROSS_SCHEDULE_ABS[operands] and cos_rows(got, want).min() >= GATE_COS
        # empty shard: empty in, empty out (embedding_generator.rs:135-137)
        corpus.wait(corpus.submit(b"", np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros((0, shape.hidden_size), np.float32)))
        with pytest.raises(pkg.B200EmbedError):
            corpus.wait(12345)  # unknown ticket
    ingest.close()


def test_corpus_reports_a_failing_shard_and_keeps_working(pkg, built, tiny_tokenizer):
    path, words = tiny_tokenizer
    small = pkg.weights.BertShape(300, 128, 1, 4, 512, 64, 2)  # vocabulary smaller than the tokenizer's: ids >= 300 are out of range
    ingest = pkg.abi.Ingest(path)
    ok_words, bad_words = words[:200], words[350:]
    good = pkg.abi.split_sentences(make_text(ok_words, [5, 9, 3, 30], 1))
    bad = pkg.abi.split_sentences(make_text(bad_words, [5, 9, 3, 30], 2))
    with pkg.abi.Corpus(small, ingest, [0], seed=3) as corpus:
        o1, o2, o3 = (np.zeros((4, 128), np.float32) for _ in range(3))
        t1, t2, t3 = corpus.submit(*good, o1), corpus.submit(*bad, o2), corpus.submit(*good, o3)
        corpus.wait(t1)
        with pytest.raises(pkg.B200EmbedError):
            corpus.wait(t2)
        corpus.wait(t3)
        assert np.array_equal(o1, o3) and np.abs(o1).max() > 0
    ingest.close()
