"""TEST INFRASTRUCTURE ONLY (never imported by the product path): Python restatement of the JSON text
serde_json::to_vec writes for shared_models::TextWithEmbeddingsMessage / QueryEmbeddingResult
(/root/reference/libs/shared_models/src/lib.rs:39-52, :66-72; published at
/root/reference/services/preprocessing_service/src/main.rs:138-142).

serde_json and ryu are crates.io dependencies that are not vendored in /root/reference (Cargo.toml:
serde_json = "1.0"), so the float layout is restated from ryu 1.x `pretty::format32` and CANNOT be pinned against
the reference binary here ("parity unpinned" for the exact text). What the tests do pin: the shortest digits
against numpy's independent shortest-repr algorithm (Dragon4, unique=True), exact f32 round trip through a JSON
parser, and the schema (field names and order) against lib.rs."""
import json

import numpy as np


def format_f32(v) -> str:
    v = np.float32(v)
    if not np.isfinite(v):
        return "null"
    sign = "-" if np.signbit(v) else ""
    v = abs(v)
    if v == 0:
        return sign + "0.0"
    s = np.format_float_scientific(v, unique=True, trim="-")  # e.g. 1.2345e-05, 3.e+00 -> "3e+00"
    mant, exp = s.split("e")
    digits = mant.replace(".", "").rstrip("0") or "0"
    kk = int(exp) + 1            # value = 0.digits x 10^kk
    n = len(digits)
    k = kk - n
    if 0 <= k and kk <= 13:
        return sign + digits + "0" * k + ".0"
    if 0 < kk <= 13:
        return sign + digits[:kk] + "." + digits[kk:]
    if -6 < kk <= 0:
        return sign + "0." + "0" * (-kk) + digits
    e = kk - 1
    return sign + digits[0] + ("." + digits[1:] if n > 1 else "") + "e" + str(e)


def _string(s: str) -> str:
    return json.dumps(s, ensure_ascii=False)  # same escapes as serde_json: \" \\ \b \f \n \r \t, \u00XX for other controls


def _vector(v) -> str:
    return "[" + ",".join(format_f32(x) for x in np.asarray(v, dtype=np.float32)) + "]"


def text_with_embeddings(original_id, source_url, sentences, embeddings, model_name, timestamp_ms) -> bytes:
    items = ",".join('{"sentence_text":%s,"embedding":%s}' % (_string(s), _vector(e)) for s, e in zip(sentences, embeddings))
    return ('{"original_id":%s,"source_url":%s,"embeddings_data":[%s],"model_name":%s,"timestamp_ms":%d}'
            % (_string(original_id), _string(source_url), items, _string(model_name), timestamp_ms)).encode("utf-8")


def query_embedding_result(request_id, embedding=None, model_name=None, error_message=None) -> bytes:
    return ('{"request_id":%s,"embedding":%s,"model_name":%s,"error_message":%s}'
            % (_string(request_id), "null" if embedding is None else _vector(embedding),
               "null" if model_name is None else _string(model_name),
               "null" if error_message is None else _string(error_message))).encode("utf-8")
