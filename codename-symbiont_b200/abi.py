"""ctypes binding of libb200_embed.so -- one Python signature per declaration in include/b200_embed.h.

This is the stub a maintainer of the reference would write in Rust (INTEGRATION.md shows that one);
it is Python here because the image has no Rust toolchain. No torch types cross this boundary:
only pointers and sizes. Device memory for the `*_device` / `b200_op_*` entry points is passed as
raw addresses (e.g. `tensor.data_ptr()`).
"""
from __future__ import annotations

import ctypes as C
import os
import re
import tempfile

import numpy as np

from . import weights as W

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_EMBED_LIB") or os.path.join(HERE, "libb200_embed.so")  # the override is for A/B timing of two builds
HEADER_PATH = os.path.join(HERE, "..", "include", "b200_embed.h")

NOALLOC_MAX_ELEMS = 1 << 20  # B200_OP_NOALLOC_MAX_ELEMS
FLAG_NORMALIZE = 1
FLAG_FP32 = 2
OPERANDS_DEFAULT, OPERANDS_BF16, OPERANDS_FP16 = 0, 1, 2  # B200_EMBED_OPERANDS_*
OPERAND_NAMES = {"default": 0, "bf16": 1, "f16": 2, "fp16": 2}
ERR_NAMES = {0: "OK", -1: "INVALID_ARG", -2: "CUDA", -3: "OOM", -4: "UNSUPPORTED_SHAPE", -5: "NO_DEVICE", -6: "IO"}
KIND_NAMES = ["lengths", "embed_ln", "gemm_qkv", "attention", "gemm_attn_out", "gemm_ffn1", "gemm_ffn2", "add_ln",
              "pool", "ffn_fused", "qkv_attention"]


class B200EmbedError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"b200_embed error {code} ({ERR_NAMES.get(code, '?')}): {message}")
        self.code = code
        self.message = message


class Config(C.Structure):
    _fields_ = [("vocab_size", C.c_int32), ("hidden_size", C.c_int32), ("num_hidden_layers", C.c_int32),
                ("num_attention_heads", C.c_int32), ("intermediate_size", C.c_int32),
                ("max_position_embeddings", C.c_int32), ("type_vocab_size", C.c_int32), ("hidden_act", C.c_int32),
                ("layer_norm_eps", C.c_float), ("max_tokens_per_launch", C.c_int32), ("operand_format", C.c_int32)]


CORPUS_MAX_GPUS = 16  # B200_CORPUS_MAX_GPUS


class CorpusStats(C.Structure):
    _fields_ = [("sentences", C.c_int64), ("tokens", C.c_int64), ("micro_batches", C.c_int64),
                ("tokenizer_threads", C.c_int32), ("gpus", C.c_int32), ("tokenize_busy_s", C.c_double), ("plan_s", C.c_double),
                ("gather_s", C.c_double), ("gpu_wait_s", C.c_double), ("scatter_s", C.c_double),
                ("per_gpu_active_s", C.c_double * CORPUS_MAX_GPUS), ("per_gpu_sentences", C.c_int64 * CORPUS_MAX_GPUS),
                ("per_gpu_tokens", C.c_int64 * CORPUS_MAX_GPUS), ("per_gpu_batches", C.c_int64 * CORPUS_MAX_GPUS)]


class KindStats(C.Structure):
    _fields_ = [("total_ms", C.c_double), ("launches", C.c_int64)]


_P = C.c_void_p
_SIGNATURES = {
    "b200_embed_create": (C.c_int32, [C.POINTER(Config), C.c_char_p, C.c_uint64, C.c_int32, C.POINTER(_P)]),
    "b200_embed_destroy": (None, [_P]),
    "b200_embed_hidden_size": (C.c_int32, [_P]),
    "b200_embed_max_positions": (C.c_int32, [_P]),
    "b200_embed_operand_format": (C.c_int32, [_P]),
    "b200_embed_last_error": (C.c_char_p, []),
    "b200_embed_forward": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_uint32, _P]),
    "b200_embed_forward_varlen": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_uint32, _P]),
    "b200_embed_submit": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_uint32, _P, C.POINTER(C.c_int64)]),
    "b200_embed_wait": (C.c_int32, [_P, C.c_int64]),
    "b200_embed_submit_varlen": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_uint32, _P, C.POINTER(C.c_int64)]),
    "b200_corpus_last_error": (C.c_char_p, []),
    "b200_corpus_create": (C.c_int32, [C.POINTER(Config), C.c_char_p, C.c_uint64, _P, C.c_int32, _P, C.c_int32, C.c_int32, C.c_int32,
                                       C.POINTER(_P)]),
    "b200_corpus_destroy": (None, [_P]),
    "b200_corpus_submit": (C.c_int32, [_P, _P, _P, _P, C.c_int64, C.c_uint32, _P, C.POINTER(C.c_int64)]),
    "b200_corpus_wait": (C.c_int32, [_P, C.c_int64]),
    "b200_corpus_stats_read": (C.c_int32, [_P, C.POINTER(CorpusStats), C.c_int32]),
    "b200_corpus_plan": (C.c_int64, [_P, C.c_int64, C.c_int32, C.c_int32, C.c_int64, _P, _P, C.c_int64]),
    "b200_corpus_plan_packed": (C.c_int64, [_P, C.c_int64, C.c_int32, C.c_int32, C.c_int64, _P, _P, C.c_int64]),
    "b200_embed_forward_device": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_uint32, _P, _P]),
    "b200_embed_host_alloc": (C.c_int32, [C.c_size_t, C.POINTER(_P)]),
    "b200_embed_host_free": (None, [_P]),
    "b200_embed_export_safetensors": (C.c_int32, [_P, C.c_char_p]),
    "b200_embed_kernel_launches": (C.c_int64, [_P]),
    "b200_embed_profile_enable": (C.c_int32, [_P, C.c_int32]),
    "b200_embed_profile_read": (C.c_int32, [_P, C.POINTER(KindStats), C.c_int32]),
    "b200_op_embed_layernorm": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P]),
    "b200_op_linear": (C.c_int32, [_P, C.c_int32, C.c_int32, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P]),
    "b200_op_ffn": (C.c_int32, [_P, C.c_int32, _P, _P, C.c_int32, _P, _P, _P]),
    "b200_op_linear_noalloc": (C.c_int32, [_P, C.c_int32, C.c_int32, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P]),
    "b200_op_attention": (C.c_int32, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "b200_op_qkv_attention": (C.c_int32, [_P, C.c_int32, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "b200_op_add_layernorm": (C.c_int32, [_P, C.c_int32, C.c_int32, _P, C.c_int32, _P, _P, C.c_int32, C.c_int32, _P, _P,
                                          _P]),
    "b200_op_mean_pool": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "b200_op_gemm_bf16": (C.c_int32, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "b200_op_gemm_f16": (C.c_int32, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P]),
    "b200_json_text_with_embeddings": (C.c_int64, [C.c_char_p, C.c_char_p, _P, _P, C.c_int32, C.c_int32, C.c_char_p,
                                                   C.c_uint64, C.c_int32, _P, C.c_int64]),
    "b200_json_text_with_embeddings_ranges": (C.c_int64, [C.c_char_p, C.c_char_p, _P, _P, _P, _P, C.c_int32, C.c_int32, C.c_char_p,
                                                          C.c_uint64, C.c_int32, _P, C.c_int64]),
    "b200_json_query_embedding_result": (C.c_int64, [C.c_char_p, _P, C.c_int32, C.c_char_p, C.c_char_p, _P, C.c_int64]),
    "b200_json_format_f32": (C.c_int32, [C.c_float, _P]),
    "b200_json_envelope": (C.c_int64, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64, _P, C.c_int64, C.POINTER(C.c_int64), _P, C.c_int64,
                                       C.POINTER(C.c_int64)]),
    "b200_json_sentence_items_ranges": (C.c_int64, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, C.c_int64]),
    "b200_jsongpu_last_error": (C.c_char_p, []),
    "b200_jsongpu_create": (C.c_int32, [C.c_int32, C.POINTER(_P)]),
    "b200_jsongpu_destroy": (None, [_P]),
    "b200_jsongpu_sentence_items": (C.c_int64, [_P, _P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, C.c_int64]),
    "b200_jsongpu_sentence_items_host": (C.c_int64, [_P, _P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, C.c_int64]),
    "b200_pipeline_last_error": (C.c_char_p, []),
    "b200_pipeline_text_to_json": (C.c_int32, [_P, _P, _P, C.c_int64, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64, C.c_int32, C.c_uint32,
                                               _P, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "b200_index_last_error": (C.c_char_p, []),
    "b200_index_create": (C.c_int32, [C.c_int32, C.c_int64, C.c_int32, C.POINTER(_P)]),
    "b200_index_destroy": (None, [_P]),
    "b200_index_size": (C.c_int64, [_P]),
    "b200_index_dim": (C.c_int32, [_P]),
    "b200_index_add": (C.c_int32, [_P, _P, C.c_int64]),
    "b200_index_add_device": (C.c_int32, [_P, _P, C.c_int64]),
    "b200_index_search": (C.c_int32, [_P, _P, C.c_int32, C.c_int32, _P, _P]),
    "b200_index_last_scan_ms": (C.c_float, [_P]),
    "b200_index_enable_prescan": (C.c_int32, [_P]),
    "b200_index_exact_fallbacks": (C.c_int64, [_P]),
    "b200_index_save": (C.c_int32, [_P, C.c_char_p]),
    "b200_index_load": (C.c_int32, [C.c_char_p, C.c_int32, C.POINTER(_P)]),
    "b200_ingest_last_error": (C.c_char_p, []),
    "b200_ingest_create": (C.c_int32, [C.c_char_p, C.POINTER(_P)]),
    "b200_ingest_destroy": (None, [_P]),
    "b200_ingest_vocab_size": (C.c_int32, [_P]),
    "b200_ingest_pad_id": (C.c_int32, [_P]),
    "b200_ingest_split_sentences": (C.c_int64, [_P, C.c_int64, _P, C.c_int64, C.POINTER(C.c_int64), _P, _P, C.c_int64]),
    "b200_ingest_split_sentences_threads": (C.c_int64, [_P, C.c_int64, _P, C.c_int64, C.POINTER(C.c_int64), _P, _P, C.c_int64,
                                                        C.c_int32]),
    "b200_ingest_encode_packed": (C.c_int64, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P, C.c_int64, _P]),
    "b200_ingest_encode": (C.c_int32, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P, _P]),
}

_lib = None


def declared_symbols() -> list:
    """Every function name include/b200_embed.h declares."""
    with open(HEADER_PATH) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_(?:embed|op|json|jsongpu|ingest|index|pipeline|corpus)_[a-z0-9_]+)\s*\(", text)))


def load_library():
    """Loads libb200_embed.so; raises if it has not been built. There is no fallback implementation."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python __graft_entry__.py` (build()) first; "
                              "there is no CPU or PyTorch fallback for this path")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def missing_symbols(lib=None) -> list:
    lib = lib or load_library()
    out = [n for n in declared_symbols() if not hasattr(lib, n)]
    out += [n for n in declared_symbols() if n not in _SIGNATURES]
    return out


def _check(rc: int):
    if rc != 0:
        raise B200EmbedError(rc, load_library().b200_embed_last_error().decode(errors="replace"))


def _u32(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    return a, a.ctypes.data


def _json_out(call, cap: int, as_array: bool):
    buf = np.empty(cap, dtype=np.uint8)  # upper bound: one formatting pass
    need = call(buf.ctypes.data, cap)
    if need < 0:
        raise ValueError("bad argument")
    if need > cap:  # cannot happen with the bounds used below; kept as the documented fallback
        buf = np.empty(need, dtype=np.uint8)
        call(buf.ctypes.data, need)
    return buf[:need] if as_array else buf[:need].tobytes()


def json_text_with_embeddings(original_id: str, source_url: str, sentences, embeddings: np.ndarray, model_name: str,
                              timestamp_ms: int, threads: int = 0, as_array: bool = False):
    """serde_json::to_vec(&TextWithEmbeddingsMessage) (libs/shared_models/src/lib.rs:45-52) for these fields.
    `sentences` is a list of str, or (text: bytes, begin, end) as split_sentences returns them."""
    lib = load_library()
    emb = np.ascontiguousarray(embeddings, dtype=np.float32)
    ranges = isinstance(sentences, tuple)
    n = len(sentences[1]) if ranges else len(sentences)
    dim = emb.shape[1] if emb.ndim == 2 else 0
    if emb.ndim != 2 or emb.shape[0] != n:
        raise ValueError("embeddings must be [len(sentences), dim]")
    tail = (emb.ctypes.data_as(_P), n, dim, model_name.encode(), int(timestamp_ms), int(threads))
    head = (original_id.encode(), source_url.encode())
    fixed = n * (dim * 17 + 48) + 6 * (len(head[0]) + len(head[1]) + len(tail[3])) + 4096
    if ranges:
        text, begin, end = sentences
        begin = np.ascontiguousarray(begin, dtype=np.int64)
        end = np.ascontiguousarray(end, dtype=np.int64)
        cap = fixed + int((end - begin).sum()) * 6
        return _json_out(lambda p, c: lib.b200_json_text_with_embeddings_ranges(*head, C.c_char_p(text), begin.ctypes.data, end.ctypes.data,
                                                                               *tail, p, c), cap, as_array)
    enc = [s.encode("utf-8") for s in sentences]
    arr = (C.c_char_p * max(n, 1))(*enc)
    cap = fixed + sum(len(e) for e in enc) * 6
    return _json_out(lambda p, c: lib.b200_json_text_with_embeddings(*head, C.cast(arr, _P), *tail, p, c), cap, as_array)


def json_query_embedding_result(request_id: str, embedding=None, model_name=None, error_message=None) -> bytes:
    """serde_json::to_vec(&QueryEmbeddingResult) (libs/shared_models/src/lib.rs:66-72); None -> null."""
    lib = load_library()
    emb = None if embedding is None else np.ascontiguousarray(embedding, dtype=np.float32)
    args = (request_id.encode(), None if emb is None else emb.ctypes.data_as(_P), 0 if emb is None else emb.size,
            None if model_name is None else model_name.encode(), None if error_message is None else error_message.encode())
    need = lib.b200_json_query_embedding_result(*args, None, 0)
    buf = C.create_string_buffer(max(need, 1))
    lib.b200_json_query_embedding_result(*args, C.cast(buf, _P), need)
    return buf.raw[:need]


def json_format_f32(v: float) -> str:
    buf = C.create_string_buffer(24)
    n = load_library().b200_json_format_f32(C.c_float(v), C.cast(buf, _P))
    return buf.raw[:n].decode()


def _check_ingest(rc: int):
    if rc < 0:
        raise B200EmbedError(int(rc), load_library().b200_ingest_last_error().decode(errors="replace"))


def split_sentences(raw_text, threads: int = 0):
    """process_raw_text_message's splitter (preprocessing_service/src/main.rs:28-63) -> (cleaned: bytes, begin, end);
    sentence i is cleaned[begin[i]:end[i]]. An empty cleaned text raises like the reference's Err("Cleaned text is empty")."""
    lib = load_library()
    raw = raw_text if isinstance(raw_text, (bytes, bytearray)) else raw_text.encode("utf-8")
    cleaned = np.empty(max(len(raw), 1), dtype=np.uint8)
    clen = C.c_int64(0)
    cap = len(raw) // 16 + 16  # a guess; the call returns the full count, so a second call fixes an underestimate
    while True:
        begin = np.empty(cap, dtype=np.int64)
        end = np.empty(cap, dtype=np.int64)
        n = lib.b200_ingest_split_sentences_threads(raw, len(raw), cleaned.ctypes.data, len(cleaned), C.byref(clen),
                                                    begin.ctypes.data, end.ctypes.data, cap, threads)
        if n <= cap:
            break
        cap = int(n)
    _check_ingest(n)
    if n == 0:
        raise ValueError("Cleaned text is empty")
    return cleaned[:clen.value].tobytes(), begin[:n], end[:n]


class Ingest:
    """Owns one b200_ingest handle: the WordPiece tokenizer of a tokenizer.json (embedding_generator.rs:73-99, :160-187)."""

    def __init__(self, tokenizer_json_path: str):
        self._h = _P()
        _check_ingest(load_library().b200_ingest_create(os.fsencode(tokenizer_json_path), C.byref(self._h)))

    def close(self):
        if self._h:
            load_library().b200_ingest_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def vocab_size(self) -> int:
        return load_library().b200_ingest_vocab_size(self._h)

    @property
    def pad_id(self) -> int:
        return load_library().b200_ingest_pad_id(self._h)

    def encode_ranges(self, text: bytes, begin, end, seq_len: int, threads: int = 0, out=None):
        """Sentence i = text[begin[i]:end[i]] (UTF-8) -> ids, mask, type_ids [n, seq_len] u32 and lengths [n] i32."""
        begin = np.ascontiguousarray(begin, dtype=np.int64)
        end = np.ascontiguousarray(end, dtype=np.int64)
        n = len(begin)
        if out is None:
            out = tuple(np.empty((n, seq_len), dtype=np.uint32) for _ in range(3))
        ids, mask, tt = out
        lengths = np.empty(n, dtype=np.int32)
        _check_ingest(load_library().b200_ingest_encode(self._h, C.c_char_p(text), begin.ctypes.data, end.ctypes.data, n, seq_len,
                                                        threads, ids.ctypes.data, mask.ctypes.data, tt.ctypes.data,
                                                        lengths.ctypes.data))
        return ids, mask, tt, lengths

    def encode_ranges_packed(self, text: bytes, begin, end, max_len: int, threads: int = 0, out=None):
        """Sentence i = text[begin[i]:end[i]] -> (ids_packed, type_ids_packed, cu_seqlens): the input of forward_varlen.
        `out` = (ids, type_ids) u32 arrays to fill (e.g. pinned); they are grown if too small."""
        begin = np.ascontiguousarray(begin, dtype=np.int64)
        end = np.ascontiguousarray(end, dtype=np.int64)
        n = len(begin)
        cu = np.empty(n + 1, dtype=np.int32)
        ids, tt = out if out is not None else (np.empty(n * 32 + 64, dtype=np.uint32), np.empty(n * 32 + 64, dtype=np.uint32))
        while True:
            cap = min(ids.size, tt.size)
            total = load_library().b200_ingest_encode_packed(self._h, C.c_char_p(text), begin.ctypes.data, end.ctypes.data, n, max_len,
                                                             threads, ids.ctypes.data, tt.ctypes.data, cap, cu.ctypes.data)
            _check_ingest(total)
            if total <= cap:
                return ids[:total], tt[:total], cu
            ids, tt = np.empty(total, dtype=np.uint32), np.empty(total, dtype=np.uint32)

    def encode(self, sentences, seq_len: int, threads: int = 0):
        """tokenizer.encode_batch(sentences, true) with Fixed(seq_len) padding and LongestFirst truncation."""
        enc = [s.encode("utf-8") for s in sentences]
        end = np.cumsum([len(b) for b in enc], dtype=np.int64) if enc else np.zeros(0, dtype=np.int64)
        begin = np.concatenate([np.zeros(1, dtype=np.int64), end[:-1]]) if enc else end
        return self.encode_ranges(b"".join(enc), begin, end, seq_len, threads)


class Index:
    """Owns one b200_index handle: exact cosine top-k over vectors resident on one GPU (vector_memory_service's Qdrant
    collection with Distance::Cosine, main.rs:34-37, :140-202, :261-286)."""

    def __init__(self, dim: int, capacity_hint: int = 0, device: int = 0):
        self._h = _P()
        self._check(load_library().b200_index_create(dim, capacity_hint, device, C.byref(self._h)))
        self.dim = dim

    @staticmethod
    def _check(rc: int):
        if rc != 0:
            raise B200EmbedError(rc, load_library().b200_index_last_error().decode(errors="replace"))

    def close(self):
        if self._h:
            load_library().b200_index_destroy(self._h)
            self._h = _P()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __len__(self) -> int:
        return int(load_library().b200_index_size(self._h))

    def add(self, vectors: np.ndarray) -> None:
        v = np.ascontiguousarray(vectors, dtype=np.float32)
        if v.ndim != 2 or v.shape[1] != self.dim:
            raise ValueError("vectors must be [n, dim]")
        self._check(load_library().b200_index_add(self._h, v.ctypes.data, v.shape[0]))

    def add_device(self, d_vectors: int, n: int) -> None:
        self._check(load_library().b200_index_add_device(self._h, d_vectors, n))

    def search(self, queries: np.ndarray, top_k: int):
        """-> (ids [q, top_k] int64, scores [q, top_k] float32), best first."""
        qv = np.ascontiguousarray(queries, dtype=np.float32)
        if qv.ndim == 1:
            qv = qv[None, :]
        if qv.ndim != 2 or qv.shape[1] != self.dim:
            raise ValueError("queries must be [q, dim]")
        ids = np.empty((qv.shape[0], top_k), dtype=np.int64)
        scores = np.empty((qv.shape[0], top_k), dtype=np.float32)
        self._check(load_library().b200_index_search(self._h, qv.ctypes.data, qv.shape[0], top_k, ids.ctypes.data, scores.ctypes.data))
        return ids, scores

    @property
    def last_scan_ms(self) -> float:
        return float(load_library().b200_index_last_scan_ms(self._h))

    def enable_prescan(self) -> None:
        """Keep a bf16 copy of the rows and search in two stages (same answers, half the scan bytes)."""
        self._check(load_library().b200_index_enable_prescan(self._h))

    @property
    def exact_fallbacks(self) -> int:
        return int(load_library().b200_index_exact_fallbacks(self._h))

    def save(self, path: str) -> None:
        self._check(load_library().b200_index_save(self._h, os.fsencode(path)))

    @classmethod
    def load(cls, path: str, device: int = 0) -> "Index":
        self = cls.__new__(cls)
        self._h = _P()
        cls._check(load_library().b200_index_load(os.fsencode(path), device, C.byref(self._h)))
        self.dim = int(load_library().b200_index_dim(self._h))
        return self


def pipeline_text_to_json(embedder, ingest, raw_text: str, original_id: str, source_url: str, model_name: str, timestamp_ms: int,
                          chunk_sentences: int = 0, flags: int = 0):
    """b200_pipeline_text_to_json: raw text -> (JSON bytes as a uint8 array, number of sentences), stages overlapped."""
    lib = load_library()
    raw = raw_text.encode("utf-8")
    out_len, n = C.c_int64(0), C.c_int64(0)
    args = (embedder.h, ingest._h, raw, len(raw), original_id.encode(), source_url.encode(), model_name.encode(), int(timestamp_ms),
            int(chunk_sentences), int(flags))
    rc = lib.b200_pipeline_text_to_json(*args, None, 0, C.byref(out_len), C.byref(n))  # learns the capacity to provide
    if out_len.value <= 0:
        raise B200EmbedError(rc, lib.b200_pipeline_last_error().decode(errors="replace"))
    buf = np.empty(out_len.value, dtype=np.uint8)
    rc = lib.b200_pipeline_text_to_json(*args, buf.ctypes.data, buf.size, C.byref(out_len), C.byref(n))
    if rc != 0:
        raise B200EmbedError(rc, lib.b200_pipeline_last_error().decode(errors="replace"))
    return buf[:out_len.value], int(n.value)


def host_alloc(shape, dtype) -> np.ndarray:
    """A numpy array in pinned host memory (b200_embed_host_alloc); free with host_free(arr)."""
    lib = load_library()
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = _P()
    _check(lib.b200_embed_host_alloc(max(n, 1), C.byref(p)))
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[arr.ctypes.data] = p.value
    return arr


_PINNED = {}


def host_free(arr: np.ndarray) -> None:
    p = _PINNED.pop(arr.ctypes.data, None)
    if p is not None:
        load_library().b200_embed_host_free(p)


class Embedder:
    """Owns one b200_embed handle (one model replica on one GPU)."""

    def __init__(self, shape: W.BertShape, safetensors_path: str | None = None, seed: int = 1234, device: int = 0,
                 max_tokens_per_launch: int = 0, operands: str | int = "default"):
        """operands: "f16" (the default: IEEE fp16 tensor-core operands) or "bf16" (cfg.operand_format)."""
        self.lib = load_library()
        self.shape = shape
        fmt = OPERAND_NAMES[operands] if isinstance(operands, str) else int(operands)
        cfg = Config(shape.vocab_size, shape.hidden_size, shape.num_hidden_layers, shape.num_attention_heads,
                     shape.intermediate_size, shape.max_position_embeddings, shape.type_vocab_size,
                     shape.hidden_act, shape.layer_norm_eps, max_tokens_per_launch, fmt)
        h = _P()
        path = safetensors_path.encode() if safetensors_path else None
        _check(self.lib.b200_embed_create(C.byref(cfg), path, seed, device, C.byref(h)))
        self.h = h
        self.device = device

    @classmethod
    def random_init(cls, shape, seed=1234, device=0, **kw):
        return cls(shape, None, seed, device, **kw)

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200_embed_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def hidden_size(self) -> int:
        return self.lib.b200_embed_hidden_size(self.h)

    @property
    def max_positions(self) -> int:
        return self.lib.b200_embed_max_positions(self.h)

    @property
    def operands(self) -> str:
        """"f16" or "bf16": the 16-bit format the handle computes in (and its b200_op_* entries take tensors in)."""
        return "f16" if self.lib.b200_embed_operand_format(self.h) == OPERANDS_FP16 else "bf16"

    def kernel_launches(self) -> int:
        return int(self.lib.b200_embed_kernel_launches(self.h))

    # -- host-buffer entry points ---------------------------------------------------------------
    def forward(self, ids, mask, type_ids, flags: int = 0, out: np.ndarray | None = None) -> np.ndarray:
        ids, p_ids = _u32(ids)
        mask, p_mask = _u32(mask)
        type_ids, p_tt = _u32(type_ids)
        if ids.ndim != 2 or ids.shape != mask.shape or ids.shape != type_ids.shape:
            raise ValueError("ids, mask and type_ids must be [batch, seq_len] arrays of one shape")
        b, s = ids.shape
        if out is None:
            out = np.empty((b, self.shape.hidden_size), dtype=np.float32)
        _check(self.lib.b200_embed_forward(self.h, p_ids, p_mask, p_tt, b, s, flags, out.ctypes.data))
        return out

    def forward_varlen(self, ids_packed, type_ids_packed, cu_seqlens, max_seqlen: int, flags: int = 0,
                       out: np.ndarray | None = None) -> np.ndarray:
        ids, p_ids = _u32(ids_packed)
        tt, p_tt = _u32(type_ids_packed)
        cu = np.ascontiguousarray(cu_seqlens, dtype=np.int32)
        b = len(cu) - 1
        if out is None:
            out = np.empty((b, self.shape.hidden_size), dtype=np.float32)
        _check(self.lib.b200_embed_forward_varlen(self.h, p_ids, p_tt, cu.ctypes.data, b, int(ids.size), max_seqlen,
                                                  flags, out.ctypes.data))
        return out

    def submit(self, ids, mask, type_ids, out: np.ndarray, flags: int = 0) -> int:
        """Buffers must stay alive (and should be pinned) until wait(ticket)."""
        for a in (ids, mask, type_ids):
            assert a.dtype == np.uint32 and a.flags.c_contiguous
        assert out.dtype == np.float32 and out.flags.c_contiguous
        b, s = ids.shape
        t = C.c_int64()
        _check(self.lib.b200_embed_submit(self.h, ids.ctypes.data, mask.ctypes.data, type_ids.ctypes.data, b, s, flags,
                                          out.ctypes.data, C.byref(t)))
        return t.value

    def wait(self, ticket: int) -> None:
        _check(self.lib.b200_embed_wait(self.h, ticket))

    def submit_varlen(self, ids_packed, type_ids_packed, cu_seqlens, max_seqlen: int, out: np.ndarray, flags: int = 0) -> int:
        """Packed form of submit: buffers must stay alive (and should be pinned) until wait(ticket)."""
        for a in (ids_packed, type_ids_packed):
            assert a.dtype == np.uint32 and a.flags.c_contiguous
        assert cu_seqlens.dtype == np.int32 and out.dtype == np.float32 and out.flags.c_contiguous
        t = C.c_int64()
        _check(self.lib.b200_embed_submit_varlen(self.h, ids_packed.ctypes.data, type_ids_packed.ctypes.data, cu_seqlens.ctypes.data,
                                                 len(cu_seqlens) - 1, int(cu_seqlens[-1]), max_seqlen, flags, out.ctypes.data,
                                                 C.byref(t)))
        return t.value

    # -- device-buffer entry point --------------------------------------------------------------
    def forward_device(self, d_ids: int, d_mask: int, d_type_ids: int, batch: int, seq_len: int, d_out: int,
                       flags: int = 0, stream: int = 0) -> None:
        _check(self.lib.b200_embed_forward_device(self.h, d_ids, d_mask, d_type_ids, batch, seq_len, flags, d_out,
                                                  stream or None))

    # -- introspection ----------------------------------------------------------------------------
    def export_safetensors(self, path: str) -> None:
        _check(self.lib.b200_embed_export_safetensors(self.h, path.encode()))

    def export_weights(self) -> dict:
        """The f32 master weights the handle computes with, as {HF name: ndarray}."""
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "model.safetensors")
            self.export_safetensors(p)
            return W.load_safetensors(p)

    def profile(self, on: bool) -> None:
        _check(self.lib.b200_embed_profile_enable(self.h, int(on)))

    def profile_read(self, reset: bool = True) -> dict:
        arr = (KindStats * len(KIND_NAMES))()
        _check(self.lib.b200_embed_profile_read(self.h, arr, int(reset)))
        return {n: {"total_ms": arr[i].total_ms, "launches": int(arr[i].launches)} for i, n in enumerate(KIND_NAMES)}

    # -- single kernels (device pointers) -------------------------------------------------------
    # every *_lo argument is the low half of a (hi, lo) activation: rows x H BYTES, one e5m2 float per element
    # (torch.float8_e5m2 has the same encoding); 0 = none
    def op_embed_layernorm(self, d_ids, d_tt, d_cu, batch, total, dtype, d_out, d_out_lo=0, stream=0):
        _check(self.lib.b200_op_embed_layernorm(self.h, d_ids, d_tt, d_cu, batch, total, dtype, d_out, d_out_lo or None,
                                                stream or None))

    def op_linear(self, layer, which, d_x, d_residual, rows, dtype, d_y, d_residual_lo=0, d_y_lo=0, stream=0):
        _check(self.lib.b200_op_linear(self.h, layer, which, d_x, d_residual or None, d_residual_lo or None, rows, dtype,
                                       d_y, d_y_lo or None, stream or None))

    def op_ffn(self, layer, d_x, d_x_lo, rows, d_y, d_y_lo=0, stream=0):
        _check(self.lib.b200_op_ffn(self.h, layer, d_x, d_x_lo or None, rows, d_y, d_y_lo or None, stream or None))

    def op_linear_noalloc(self, layer, which, d_x, d_residual, rows, dtype, d_y, d_residual_lo=0, d_y_lo=0, stream=0):
        """b200_op_linear for which = 1 / 3 with the fp32 temporary in the library's built-in 4 MiB scratch: no
        cudaMalloc / cudaFree. rows * hidden must fit NOALLOC_MAX_ELEMS; overlapping calls race on the scratch."""
        _check(self.lib.b200_op_linear_noalloc(self.h, layer, which, d_x, d_residual or None, d_residual_lo or None, rows,
                                               dtype, d_y, d_y_lo or None, stream or None))

    def op_attention(self, d_qkv, d_cu, batch, rows, max_seqlen, dtype, d_ctx, stream=0):
        _check(self.lib.b200_op_attention(self.h, d_qkv, d_cu, batch, rows, max_seqlen, dtype, d_ctx, stream or None))

    def op_qkv_attention(self, layer, d_x, d_cu, batch, rows, max_seqlen, d_ctx, stream=0):
        _check(self.lib.b200_op_qkv_attention(self.h, layer, d_x, d_cu, batch, rows, max_seqlen, d_ctx, stream or None))

    def op_add_layernorm(self, layer, which, d_y, y_is_f32, d_residual, rows, dtype, d_out, d_residual_lo=0,
                         d_out_lo=0, stream=0):
        _check(self.lib.b200_op_add_layernorm(self.h, layer, which, d_y, int(y_is_f32), d_residual, d_residual_lo or None,
                                              rows, dtype, d_out, d_out_lo or None, stream or None))

    def op_mean_pool(self, d_hidden, d_cu, batch, dtype, normalize, d_out, d_hidden_lo=0, stream=0):
        _check(self.lib.b200_op_mean_pool(self.h, d_hidden, d_hidden_lo or None, d_cu, batch, dtype, int(normalize), d_out,
                                          stream or None))


def json_sentence_items_ranges(text: bytes, begin, end, embeddings: np.ndarray, first: bool = True, threads: int = 0) -> np.ndarray:
    """b200_json_sentence_items_ranges (host): the comma-separated items of the sentences text[begin[i]:end[i]]."""
    lib = load_library()
    begin = np.ascontiguousarray(begin, dtype=np.int64)
    end = np.ascontiguousarray(end, dtype=np.int64)
    emb = np.ascontiguousarray(embeddings, dtype=np.float32)
    n, dim = emb.shape
    cap = n * (dim * 17 + 80) + int((end - begin).sum()) * 6 + 64
    out = np.empty(cap, dtype=np.uint8)
    got = lib.b200_json_sentence_items_ranges(C.c_char_p(text), begin.ctypes.data, end.ctypes.data, emb.ctypes.data, n, dim, int(first),
                                              threads, out.ctypes.data, cap)
    if got < 0:
        raise B200EmbedError(int(got), "b200_json_sentence_items_ranges failed")
    return out[:got]


class JsonGpu:
    """b200_jsongpu_*: the JSON items of sentences formatted on the GPU from device-resident embeddings."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = _P()
        rc = self.lib.b200_jsongpu_create(device, C.byref(h))
        if rc != 0:
            raise B200EmbedError(rc, self.lib.b200_jsongpu_last_error().decode())
        self.h = h

    def sentence_items(self, text: bytes, begin, end, d_embeddings: int, n: int, dim: int, first: bool = True, out: np.ndarray | None = None):
        """-> uint8 array holding the items (a view of `out` when given and large enough; `out` should be pinned)."""
        begin = np.ascontiguousarray(begin, dtype=np.int64)
        end = np.ascontiguousarray(end, dtype=np.int64)
        cap = n * (dim * 17 + 48) + int((end - begin).sum()) * 6 + 64
        if out is None or out.size < cap:
            out = np.empty(cap, dtype=np.uint8)
        got = self.lib.b200_jsongpu_sentence_items(self.h, C.c_char_p(text), begin.ctypes.data, end.ctypes.data, d_embeddings, n, dim,
                                                   int(first), out.ctypes.data, out.size)
        if got < 0:
            raise B200EmbedError(int(got), self.lib.b200_jsongpu_last_error().decode())
        return out[:got]

    def sentence_items_host(self, text: bytes, begin, end, embeddings: np.ndarray, first: bool = True, out: np.ndarray | None = None):
        """The same from host embeddings ([n, dim] f32, pinned for speed): uploaded, formatted, the text copied back."""
        begin = np.ascontiguousarray(begin, dtype=np.int64)
        end = np.ascontiguousarray(end, dtype=np.int64)
        assert embeddings.dtype == np.float32 and embeddings.flags.c_contiguous
        n, dim = embeddings.shape
        cap = n * (dim * 17 + 48) + int((end - begin).sum()) * 6 + 64
        if out is None or out.size < cap:
            out = np.empty(cap, dtype=np.uint8)
        got = self.lib.b200_jsongpu_sentence_items_host(self.h, C.c_char_p(text), begin.ctypes.data, end.ctypes.data, embeddings.ctypes.data,
                                                        n, dim, int(first), out.ctypes.data, out.size)
        if got < 0:
            raise B200EmbedError(int(got), self.lib.b200_jsongpu_last_error().decode())
        return out[:got]

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200_jsongpu_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def corpus_plan(lengths, max_len: int, max_batch_sentences: int, max_batch_tokens: int):
    """b200_corpus_plan: -> (order [n] int32: the sentences by ascending token count, batch_begin [nb + 1] int64)."""
    lengths = np.ascontiguousarray(lengths, dtype=np.int32)
    n = len(lengths)
    order = np.empty(n, dtype=np.int32)
    begins = np.empty(n + 2, dtype=np.int64)
    nb = load_library().b200_corpus_plan(lengths.ctypes.data, n, max_len, max_batch_sentences, max_batch_tokens,
                                         order.ctypes.data, begins.ctypes.data, len(begins))
    if nb < 0:
        raise B200EmbedError(int(nb), load_library().b200_corpus_last_error().decode())
    return order, begins[:nb + 1].copy()


def corpus_plan_packed(lengths, max_len: int, max_batch_sentences: int, max_batch_tokens: int):
    """b200_corpus_plan_packed, the plan b200_corpus_submit makes: -> (order [n] int32, batch_begin [nb + 1] int64); short
    sentences in input-order micro-batches, tile by tile inside each."""
    lengths = np.ascontiguousarray(lengths, dtype=np.int32)
    n = len(lengths)
    order = np.empty(n, dtype=np.int32)
    begins = np.empty(n + 2, dtype=np.int64)
    nb = load_library().b200_corpus_plan_packed(lengths.ctypes.data, n, max_len, max_batch_sentences, max_batch_tokens,
                                                order.ctypes.data, begins.ctypes.data, len(begins))
    if nb < 0:
        raise B200EmbedError(int(nb), load_library().b200_corpus_last_error().decode())
    return order, begins[:nb + 1].copy()


class Corpus:
    """Owns one b200_corpus handle: shards of sentences in, embeddings out in input order, over all the listed GPUs."""

    def __init__(self, shape: W.BertShape, ingest: "Ingest", devices, safetensors_path: str | None = None, seed: int = 1234,
                 tokenizer_threads: int = 0, max_batch_sentences: int = 0, max_batch_tokens: int = 0, operands: str = "default"):
        self.lib = load_library()
        self.shape, self.ingest = shape, ingest  # the tokenizer must outlive the corpus handle
        cfg = Config(shape.vocab_size, shape.hidden_size, shape.num_hidden_layers, shape.num_attention_heads,
                     shape.intermediate_size, shape.max_position_embeddings, shape.type_vocab_size,
                     shape.hidden_act, shape.layer_norm_eps, 0, OPERAND_NAMES[operands])
        dev = np.ascontiguousarray(list(devices), dtype=np.int32)
        h = _P()
        rc = self.lib.b200_corpus_create(C.byref(cfg), safetensors_path.encode() if safetensors_path else None, seed, dev.ctypes.data,
                                         len(dev), ingest._h, tokenizer_threads, max_batch_sentences, max_batch_tokens, C.byref(h))
        if rc != 0:
            raise B200EmbedError(rc, self.lib.b200_corpus_last_error().decode())
        self.h, self._keep = h, {}

    def submit(self, text: bytes, begin, end, out: np.ndarray, flags: int = 0) -> int:
        """Shard = sentences text[begin[i]:end[i]]; out [n, hidden] f32. Everything is kept alive until wait(ticket)."""
        begin = np.ascontiguousarray(begin, dtype=np.int64)
        end = np.ascontiguousarray(end, dtype=np.int64)
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.shape == (len(begin), self.shape.hidden_size)
        t = C.c_int64()
        rc = self.lib.b200_corpus_submit(self.h, C.c_char_p(text), begin.ctypes.data, end.ctypes.data, len(begin), flags,
                                         out.ctypes.data, C.byref(t))
        if rc != 0:
            raise B200EmbedError(rc, self.lib.b200_corpus_last_error().decode())
        self._keep[t.value] = (text, begin, end, out)
        return t.value

    def wait(self, ticket: int) -> np.ndarray:
        rc = self.lib.b200_corpus_wait(self.h, ticket)
        kept = self._keep.pop(ticket, None)
        if rc != 0:
            raise B200EmbedError(rc, self.lib.b200_corpus_last_error().decode())
        return kept[3] if kept else None

    def stats(self, reset: bool = False) -> dict:
        st = CorpusStats()
        self.lib.b200_corpus_stats_read(self.h, C.byref(st), int(reset))
        g = st.gpus
        d = {k: getattr(st, k) for k in ("sentences", "tokens", "micro_batches", "tokenizer_threads", "gpus", "tokenize_busy_s",
                                         "plan_s", "gather_s", "gpu_wait_s", "scatter_s")}
        for k in ("per_gpu_active_s", "per_gpu_sentences", "per_gpu_tokens", "per_gpu_batches"):
            d[k] = list(getattr(st, k))[:g]
        return d

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200_corpus_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def op_gemm_bf16(d_a, d_b, d_bias, m, n, k, act, block_n, d_c, stream=0):
    _check(load_library().b200_op_gemm_bf16(d_a, d_b, d_bias, m, n, k, act, block_n, d_c, stream or None))


def op_gemm_f16(d_a, d_b, d_bias, m, n, k, act, block_n, d_c, stream=0):
    _check(load_library().b200_op_gemm_f16(d_a, d_b, d_bias, m, n, k, act, block_n, d_c, stream or None))
