"""Host-side mirror of the reference's `EmbeddingGenerator`
(/root/reference/services/preprocessing_service/src/embedding_generator.rs:9-224): same two entry
points, same argument meaning, same error behaviour, with the tensor work behind the C ABI.

    EmbeddingGenerator.new(model_dir, revision, force_cpu)          (:17)
    generator.generate_sentence_embeddings(sentences) -> [[f32]]    (:134)

What is unchanged from the reference: the tokenizer's behaviour (a BERT WordPiece tokenizer.json, and the Unigram /
SentencePiece tokenizer.json of XLM-R checkpoints such as the service's deployed model, go through the native
b200_ingest_* entries, checked bit-exactly against the `tokenizers` library in tests/test_ingest.py and
tests/test_ingest_unigram.py; any other tokenizer.json goes through the `tokenizers` library itself, the Python binding
of the Rust core the reference links), Fixed right padding to `max_position_embeddings` with `[PAD]`
(:75-91), LongestFirst truncation to the same length (:93-99), `encode_batch(inputs, true)` (:160-164),
the length check (:172-177), the flattening of ids / attention mask / type ids (:179-187), the order of
the results and "empty in => empty out" (:135-137). What changes: lines :189-214 (three candle tensors,
`BertModel::forward`, the pooling ops, `to_vec2`) become one `b200_embed_forward` call, and the
chunk-of-8 loop (:146-148) becomes one call for all sentences -- the library never computes padded
positions, so batching cannot change a result.

Differences that are deliberate and documented: there is no network here, so `model_id` is a
local directory holding `tokenizer.json`, `config.json` and `model.safetensors` instead of a hub id (:25-58);
`force_cpu=True` raises, because this path has no CPU fallback.
"""
from __future__ import annotations

import os
from typing import List, Sequence

import numpy as np

from . import weights as W
from .abi import Embedder, B200EmbedError, FLAG_NORMALIZE, Ingest, split_sentences

ERR_UNSUPPORTED = -4  # B200_EMBED_ERR_UNSUPPORTED_SHAPE


class EmbeddingGenerator:
    def __init__(self, embedder: Embedder, tokenizer, shape: W.BertShape, normalize: bool = False, ingest: Ingest | None = None):
        self.model = embedder
        self.tokenizer = tokenizer  # the `tokenizers` library object, used only when `ingest` is None
        self.ingest = ingest        # native front-end (b200_ingest_*: WordPiece or Unigram), None for any other tokenizer.json
        self.config = shape
        self.normalize = normalize

    @classmethod
    def new(cls, model_id: str, revision: str | None = None, force_cpu: bool = False, device: int = 0,
            normalize: bool = False, operands: str = "default") -> "EmbeddingGenerator":
        if force_cpu:
            raise B200EmbedError(-5, "force_cpu=true: this embedding path has no CPU fallback (NO_DEVICE)")
        d = model_id
        shape = W.BertShape.from_config_json(os.path.join(d, "config.json"))
        tok, ingest = None, None
        try:
            ingest = Ingest(os.path.join(d, "tokenizer.json"))
        except B200EmbedError as e:
            if e.code != ERR_UNSUPPORTED:
                raise
            from tokenizers import Tokenizer  # e.g. a BPE tokenizer

            tok = Tokenizer.from_file(os.path.join(d, "tokenizer.json"))
            configure_tokenizer(tok, shape.max_position_embeddings)
        emb = Embedder(shape, os.path.join(d, "model.safetensors"), device=device, operands=operands)
        return cls(emb, tok, shape, normalize, ingest)

    def generate_sentence_embeddings(self, sentences: Sequence[str]) -> List[List[float]]:
        if len(sentences) == 0:
            return []
        return self.generate_sentence_embeddings_array(sentences).tolist()

    def generate_sentence_embeddings_array(self, sentences: Sequence[str]) -> np.ndarray:
        """Same call, returning the [n, hidden] float32 matrix without the nested-list conversion."""
        if len(sentences) == 0:
            return np.empty((0, self.config.hidden_size), dtype=np.float32)
        P = self.config.max_position_embeddings
        if self.ingest is not None:
            ids, mask, type_ids, _ = self.ingest.encode(sentences, P)
        else:
            ids, mask, type_ids = tokenize_fixed(self.tokenizer, sentences, P)
        flags = FLAG_NORMALIZE if self.normalize else 0
        return self.model.forward(ids, mask, type_ids, flags=flags)

    def process_raw_text(self, raw_text: str, doc_id: str = ""):
        """process_raw_text_message up to the embeddings (preprocessing_service/src/main.rs:28-95): clean, split into
        sentences, embed. Returns (sentences, [n, hidden] float32); raises ValueError with the reference's messages."""
        try:
            cleaned, begin, end = split_sentences(raw_text)
        except ValueError:
            raise ValueError(f"Cleaned text is empty for id: {doc_id}") from None
        if self.ingest is None:
            sentences = [cleaned[b:e].decode() for b, e in zip(begin, end)]
            return sentences, self.generate_sentence_embeddings_array(sentences)
        # packed tokens (no padding materialised on the host or copied to the device) -> b200_embed_forward_varlen
        P = self.config.max_position_embeddings
        ids, type_ids, cu = self.ingest.encode_ranges_packed(cleaned, begin, end, P)
        flags = FLAG_NORMALIZE if self.normalize else 0
        out = self.model.forward_varlen(ids, type_ids, cu, int(np.diff(cu).max()), flags=flags)
        return [cleaned[b:e].decode() for b, e in zip(begin, end)], out

    def close(self):
        self.model.close()


def configure_tokenizer(tok, max_len: int) -> None:
    """embedding_generator.rs:75-99: pad id from `[PAD]` (0 if absent), Fixed right padding, LongestFirst truncation."""
    pad_id = tok.token_to_id("[PAD]")
    if pad_id is None:
        pad_id = 0
    tok.enable_padding(direction="right", pad_id=pad_id, pad_type_id=0, pad_token="[PAD]", length=max_len)
    tok.enable_truncation(max_length=max_len, stride=0, strategy="longest_first", direction="right")


def tokenize_fixed(tok, sentences: Sequence[str], max_len: int):
    """embedding_generator.rs:160-187: encode_batch(inputs, add_special_tokens=true) -> three [n, max_len] u32 arrays."""
    enc = tok.encode_batch(list(sentences), add_special_tokens=True)
    if not enc:
        raise RuntimeError("Empty encodings for a non-empty sentence batch, this should not happen.")
    if len(enc[0].ids) != max_len:
        raise RuntimeError(f"Tokenizer returned sequence length {len(enc[0].ids)} but model/padding is configured for {max_len}")
    ids = np.array([e.ids for e in enc], dtype=np.uint32)
    mask = np.array([e.attention_mask for e in enc], dtype=np.uint32)
    tt = np.array([e.type_ids for e in enc], dtype=np.uint32)
    return ids, mask, tt
