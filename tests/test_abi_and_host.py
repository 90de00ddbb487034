"""CPU: the C-ABI library loads and exports what include/b200_embed.h declares; host-side logic."""
import json
import os

import numpy as np
import pytest


def test_library_exports_every_declared_symbol(pkg, built):
    names = pkg.abi.declared_symbols()
    assert len(names) >= 20 and "b200_embed_forward" in names
    assert pkg.abi.missing_symbols(built) == []


def test_no_cpu_fallback(pkg, built):
    """Without a CUDA device the product path must fail loudly (NO_DEVICE), never compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.B200EmbedError) as e:
        pkg.Embedder.random_init(pkg.weights.TINY)
    assert e.value.code == -5


def test_force_cpu_is_an_error(pkg, built, tmp_path):
    with pytest.raises(pkg.B200EmbedError):
        pkg.EmbeddingGenerator.new(str(tmp_path), None, True)


def test_product_never_imports_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dp, _, files in os.walk(os.path.join(root, "codename-symbiont_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dp, f)).read()
                for needle in ("import oracle", "from oracle", "liboracle", "oracle/", "oracle."):
                    assert needle not in text, f"{f} references the test oracle ({needle!r})"


def test_init_spec_is_deterministic_and_normal(pkg):
    a = pkg.weights.init_tensor("encoder.layer.0.output.dense.weight", (64, 256), 1234)
    b = pkg.weights.init_tensor("encoder.layer.0.output.dense.weight", (64, 256), 1234)
    c = pkg.weights.init_tensor("encoder.layer.0.output.dense.weight", (64, 256), 1235)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert abs(a.std() - 0.02) < 1e-3 and abs(a.mean()) < 1e-3
    g = pkg.weights.init_tensor("embeddings.LayerNorm.weight", (128,), 1)
    assert abs(g.mean() - 1.0) < 0.01
    # known answers for the integer part of the spec (same constants as csrc/weights_io.cpp)
    assert pkg.weights._fnv1a64("embeddings.LayerNorm.bias") == 0x597E4FE67692B8AD
    assert int(pkg.weights._splitmix64(np.array([0], dtype=np.uint64))[0]) == 0xE220A8397B1DCDAF
    kat = pkg.weights.init_tensor("embeddings.LayerNorm.bias", (4,), 1234)
    assert np.allclose(kat, [-0.03523624, -0.01201622, 0.00132841, 0.00389976], rtol=0, atol=1e-8)


def test_safetensors_roundtrip(pkg, tmp_path):
    w = pkg.weights.init_weights(pkg.weights.TINY, 7)
    p = str(tmp_path / "m.safetensors")
    pkg.weights.save_safetensors(p, w)
    r = pkg.weights.load_safetensors(p)
    assert set(r) == set(w) and all(np.array_equal(r[k], w[k]) for k in w)
    from safetensors.numpy import load_file  # the reference's loader format (safetensors crate)
    r2 = load_file(p)
    assert all(np.array_equal(r2[k], w[k]) for k in w)


def test_synth_batch_layout(pkg):
    ids, mask, tt = pkg.synth.padded_batch(30522, [2, 5, 8], 8, seed=1)
    assert ids.dtype == np.uint32 and ids.shape == (3, 8)
    assert (ids[:, 0] == 101).all() and ids[0, 1] == 102 and ids[1, 4] == 102 and ids[2, 7] == 102
    assert (ids[mask == 0] == 0).all() and (mask.sum(1) == [2, 5, 8]).all() and (tt == 0).all()
    pi, pt, cu = pkg.synth.pack(ids, mask, tt)
    assert cu.tolist() == [0, 2, 7, 15] and pi.shape == (15,) and pi[2] == 101


def _wordpiece_tokenizer(tmp_path, vocab_words):
    from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, processors
    vocab = {w: i for i, w in enumerate(["[PAD]", "[UNK]", "[CLS]", "[SEP]"] + vocab_words)}
    tok = Tokenizer(models.WordPiece(vocab, unk_token="[UNK]"))
    tok.normalizer = normalizers.BertNormalizer(lowercase=True)
    tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    tok.post_processor = processors.TemplateProcessing(single="[CLS] $A [SEP]", pair="[CLS] $A [SEP] $B:1 [SEP]:1",
                                                       special_tokens=[("[CLS]", 2), ("[SEP]", 3)])
    return tok


def test_tokenize_fixed_matches_reference_settings(pkg, tmp_path):
    """embedding_generator.rs:75-99, :160-187: Fixed right padding to P, LongestFirst truncation, specials added."""
    words = ["the", "cat", "sat", "on", "mat", ".", "##s", "dog"]
    tok = _wordpiece_tokenizer(tmp_path, words)
    from codename_symbiont_b200.embedding_generator import configure_tokenizer, tokenize_fixed
    configure_tokenizer(tok, 16)
    ids, mask, tt = tokenize_fixed(tok, ["The cat sat.", "dogs " * 40], 16)
    assert ids.shape == (2, 16) and ids.dtype == np.uint32
    assert ids[0, 0] == 2 and ids[0, mask[0].sum() - 1] == 3 and (ids[0, mask[0] == 0] == 0).all()
    assert mask[1].sum() == 16 and ids[1, 15] == 3  # truncated to P with [SEP] kept
    assert all((m[:-1] >= m[1:]).all() for m in mask)  # right-padded prefix


def test_embedding_schema_roundtrip():
    """libs/shared_models/src/lib.rs:39-52, :66-72 and its tests :193-228, :255-267: f32 vectors survive JSON."""
    emb = np.array([0.1, -0.25, 3.4028235e38, 1e-9], dtype=np.float32)
    msg = {"original_id": "id", "source_url": "u", "model_name": "m", "timestamp_ms": 1,
           "embeddings_data": [{"sentence_text": "s", "embedding": [float(x) for x in emb]}]}
    back = json.loads(json.dumps(msg))
    assert np.array_equal(np.array(back["embeddings_data"][0]["embedding"], dtype=np.float32), emb)
    res = {"request_id": "r", "embedding": [float(x) for x in emb], "model_name": None, "error_message": None}
    assert np.array_equal(np.array(json.loads(json.dumps(res))["embedding"], dtype=np.float32), emb)
