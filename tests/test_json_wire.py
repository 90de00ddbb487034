"""CPU: the JSON wire format of the results (include/b200_embed.h b200_json_*) against the Python
restatement in oracle/json_oracle.py, exact f32 round trip, and the schema of libs/shared_models/src/lib.rs."""
import json
import os
import time

import numpy as np
import pytest

from oracle import json_oracle


@pytest.fixture(scope="module")
def abi(pkg, built):
    return pkg.abi


def test_f32_layout_and_round_trip(abi):
    rng = np.random.default_rng(5)
    bits = rng.integers(0, 2**32, size=20000, dtype=np.uint64).astype(np.uint32)
    vals = np.concatenate([
        bits.view(np.float32),                                   # the whole range, subnormals, inf, nan
        rng.standard_normal(20000).astype(np.float32) * 0.1,    # embedding-like
        np.float32([0.0, -0.0, 1.0, -1.0, 0.1, 1e-5, 1e-6, 1e-7, 9.999999e12, 1e13, 1e12, 16777216.0, 3.4028235e38,
                    1e-45, 1.17549435e-38, 123456.79, 0.001234, 12.34, 1e21]),
    ])
    for v in vals:
        got = abi.json_format_f32(float(v))
        assert got == json_oracle.format_f32(v), (v, got)
        if np.isfinite(v):
            back = np.float32(json.loads(got))
            assert back.view(np.uint32) == np.float32(v).view(np.uint32) or (back == 0 and v == 0), (v, got)
            assert "e+" not in got and not got.startswith(".") and got == got.strip()


def test_text_with_embeddings_message(abi):
    rng = np.random.default_rng(1)
    sentences = ['Plain sentence.', 'Quotes " and \\ backslash?', "tabs\tnew\nlines\r!", "Ünïcödé – 文字 🙂.", "ctl\x01\x1f."]
    emb = rng.standard_normal((len(sentences), 384)).astype(np.float32)
    emb[0, :4] = [0.0, -0.0, 1e-7, 1e13]
    got = abi.json_text_with_embeddings("id-1", "https://example.org/a?b=c", sentences, emb,
                                        "sentence-transformers/paraphrase-multilingual-mpnet-base-v2", 1717171717171, threads=3)
    want = json_oracle.text_with_embeddings("id-1", "https://example.org/a?b=c", sentences, emb,
                                            "sentence-transformers/paraphrase-multilingual-mpnet-base-v2", 1717171717171)
    assert got == want
    doc = json.loads(got)
    # field names and order of TextWithEmbeddingsMessage / SentenceEmbedding (lib.rs:39-52)
    assert list(doc) == ["original_id", "source_url", "embeddings_data", "model_name", "timestamp_ms"]
    assert all(list(d) == ["sentence_text", "embedding"] for d in doc["embeddings_data"])
    assert [d["sentence_text"] for d in doc["embeddings_data"]] == sentences
    back = np.array([d["embedding"] for d in doc["embeddings_data"]], dtype=np.float32)
    assert np.array_equal(back.view(np.uint32) & 0x7fffffff, emb.view(np.uint32) & 0x7fffffff)  # exact (sign of zero aside)
    assert abi.json_text_with_embeddings("x", "", [], np.zeros((0, 384), np.float32), "m", 0) == \
        b'{"original_id":"x","source_url":"","embeddings_data":[],"model_name":"m","timestamp_ms":0}'
    # thread count must not change the bytes
    assert got == abi.json_text_with_embeddings("id-1", "https://example.org/a?b=c", sentences, emb,
                                                "sentence-transformers/paraphrase-multilingual-mpnet-base-v2", 1717171717171, threads=1)


def test_query_embedding_result(abi):
    e = np.linspace(-1, 1, 768, dtype=np.float32)
    got = abi.json_query_embedding_result("req-7", e, "model", None)
    assert got == json_oracle.query_embedding_result("req-7", e, "model", None)
    doc = json.loads(got)
    assert list(doc) == ["request_id", "embedding", "model_name", "error_message"] and doc["error_message"] is None
    assert np.array_equal(np.float32(doc["embedding"]), e)
    err = abi.json_query_embedding_result("req-8", None, None, 'failed: "x"')
    assert err == b'{"request_id":"req-8","embedding":null,"model_name":null,"error_message":"failed: \\"x\\""}'


def test_formatter_throughput_is_reported(abi, capsys):
    """Not a gate: prints MB/s of the native formatter next to Python's json.dumps on one ingest-sized message."""
    rng = np.random.default_rng(2)
    n, dim = 2048, 384
    emb = rng.standard_normal((n, dim)).astype(np.float32) * 0.05
    sentences = ["Sentence number %d of the document." % i for i in range(n)]
    t0 = time.perf_counter(); a = abi.json_text_with_embeddings("i", "u", sentences, emb, "m", 1); t1 = time.perf_counter()
    b = json.dumps({"original_id": "i", "source_url": "u",
                    "embeddings_data": [{"sentence_text": s, "embedding": e.tolist()} for s, e in zip(sentences, emb)],
                    "model_name": "m", "timestamp_ms": 1}); t2 = time.perf_counter()
    with capsys.disabled():
        print(f"\n[json] native {len(a) / 1e6 / (t1 - t0):.0f} MB/s ({n / (t1 - t0):,.0f} sentences/s), "
              f"python json.dumps {len(b) / 1e6 / (t2 - t1):.0f} MB/s")
    assert len(a) > n * dim * 4


def test_ranges_entry_writes_the_same_document(abi):
    """b200_json_text_with_embeddings_ranges: sentences as byte ranges of one text (what the splitter returns), including a
    sentence with an embedded NUL, which the char* form cannot carry."""
    rng = np.random.default_rng(2)
    sentences = ["First one.", 'Second "quoted" one?', "third without stop", "x" * 300 + "!"] * 25
    emb = rng.standard_normal((len(sentences), 96)).astype(np.float32)
    enc = [s.encode() for s in sentences]
    text = b" ".join(enc)
    begin = np.cumsum([0] + [len(e) + 1 for e in enc[:-1]])
    end = begin + np.array([len(e) for e in enc])
    a = abi.json_text_with_embeddings("i", "u", sentences, emb, "m", 7, threads=4)
    b = abi.json_text_with_embeddings("i", "u", (text, begin, end), emb, "m", 7, threads=3)
    c = abi.json_text_with_embeddings("i", "u", (text, begin, end), emb, "m", 7, threads=1, as_array=True)
    assert a == b == c.tobytes() == json_oracle.text_with_embeddings("i", "u", sentences, emb, "m", 7)
    nul = b"a\x00b."
    got = abi.json_text_with_embeddings("i", "u", (nul, np.array([0]), np.array([4])), emb[:1], "m", 7)
    assert json.loads(got)["embeddings_data"][0]["sentence_text"] == "a\x00b."


def test_f32_digit_generator_matches_the_cxx_library(tmp_path):
    """The Schubfach digit generator of csrc/json_out.cpp against std::to_chars (shortest round-trip) on every 61st
    positive finite float (35 M values, ~2 s); scripts/probes/f32_digits_check.cpp with stride 1 is the exhaustive run
    (2 139 095 039 values, 0 mismatches, recorded in DESIGN.md)."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "f32check")
    subprocess.run(["g++", "-std=c++17", "-O2", os.path.join(root, "scripts", "probes", "f32_digits_check.cpp"), "-o", exe, "-lpthread"],
                   check=True, capture_output=True)
    r = subprocess.run([exe, "61"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and " 0 mismatches" in r.stdout, r.stdout


def test_envelope_and_chunked_items_make_the_same_document(abi):
    """b200_json_envelope + b200_json_sentence_items_ranges (chunked formatting / publishing): head + items of the chunks +
    tail is byte for byte the document the one-call entry writes."""
    import ctypes as C
    lib = abi.load_library()
    rng = np.random.default_rng(3)
    sentences = ['One "quoted".', "Two\ttabs?", "three", "Four!"] * 11
    emb = rng.standard_normal((len(sentences), 64)).astype(np.float32)
    enc = [s.encode() for s in sentences]
    text = b"|".join(enc)
    begin = np.cumsum([0] + [len(e) + 1 for e in enc[:-1]]).astype(np.int64)
    end = begin + np.array([len(e) for e in enc], dtype=np.int64)
    whole = abi.json_text_with_embeddings("id", "http://u", (text, begin, end), emb, "model", 99)
    hl, tl = C.c_int64(0), C.c_int64(0)
    lib.b200_json_envelope(b"id", b"http://u", b"model", 99, None, 0, C.byref(hl), None, 0, C.byref(tl))
    head, tail = C.create_string_buffer(hl.value), C.create_string_buffer(tl.value)
    lib.b200_json_envelope(b"id", b"http://u", b"model", 99, C.cast(head, C.c_void_p), hl.value, C.byref(hl), C.cast(tail, C.c_void_p), tl.value, C.byref(tl))
    doc = head.raw[:hl.value]
    for lo, hi in ((0, 17), (17, 18), (18, len(sentences))):
        b, e, v = np.ascontiguousarray(begin[lo:hi]), np.ascontiguousarray(end[lo:hi]), np.ascontiguousarray(emb[lo:hi])
        need = lib.b200_json_sentence_items_ranges(C.c_char_p(text), b.ctypes.data, e.ctypes.data, v.ctypes.data, hi - lo, 64, int(lo == 0), 2, None, 0)
        buf = C.create_string_buffer(need)
        got = lib.b200_json_sentence_items_ranges(C.c_char_p(text), b.ctypes.data, e.ctypes.data, v.ctypes.data, hi - lo, 64, int(lo == 0), 3,
                                                  C.cast(buf, C.c_void_p), need)
        assert got == need
        doc += buf.raw[:need]
    assert doc + tail.raw[:tl.value] == whole
