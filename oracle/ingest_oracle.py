"""TEST INFRASTRUCTURE ONLY - never imported by the product path.

CPU restatement of the sentence splitter of process_raw_text_message
(/root/reference/services/preprocessing_service/src/main.rs:28-63) in plain Python, used by tests/test_ingest.py to
check b200_ingest_split_sentences. Rust's str::split_whitespace and str::trim use the Unicode White_Space property
(char::is_whitespace), which is NOT Python's str.split()/str.strip() set (Python also splits on U+001C..U+001F), so the
set is spelled out. The reference's own tests hold no fixture for the splitter: the restatement is pinned only by
reading the source ("parity unpinned" for the splitter); the tokenizer half of the ingest path is checked against the
`tokenizers` library itself, the reference's dependency.
"""
RUST_WHITESPACE = frozenset(
    [0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x20, 0x85, 0xA0, 0x1680, 0x2028, 0x2029, 0x202F, 0x205F, 0x3000] + list(range(0x2000, 0x200B)))


def _is_ws(ch: str) -> bool:
    return ord(ch) in RUST_WHITESPACE


def _trim(s: str) -> str:
    b, e = 0, len(s)
    while b < e and _is_ws(s[b]):
        b += 1
    while e > b and _is_ws(s[e - 1]):
        e -= 1
    return s[b:e]


def clean_text(raw: str) -> str:
    """main.rs:28-32: raw_text.split_whitespace().collect::<Vec<&str>>().join(" ")"""
    words, cur = [], []
    for ch in raw:
        if _is_ws(ch):
            if cur:
                words.append("".join(cur))
                cur = []
        else:
            cur.append(ch)
    if cur:
        words.append("".join(cur))
    return " ".join(words)


def split_sentences(raw: str):
    """main.rs:28-63 -> (cleaned, [sentences]); an empty cleaned text returns ("", []) (the reference returns Err)."""
    cleaned = clean_text(raw)
    if not cleaned:
        return cleaned, []
    out, start = [], 0
    for i, ch in enumerate(cleaned):  # main.rs:42-50
        if ch in ".?!":
            out.append(_trim(cleaned[start:i + 1]))
            start = i + 1
    if start < len(cleaned):  # main.rs:52-57
        rest = _trim(cleaned[start:])
        if rest:
            out.append(rest)
    if not out:  # main.rs:59-61
        out.append(cleaned)
    return cleaned, out
