"""GPU: b200_jsongpu_sentence_items, the JSON items formatted on the device, must be byte for byte what the host formatter
b200_json_sentence_items_ranges writes (which tests/test_json_wire.py pins: shortest round-trip digits, ryu's layout,
serde_json's escapes; /root/reference/libs/shared_models/src/lib.rs:39-43, preprocessing_service/src/main.rs:138)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ranges(sentences):
    enc = [s.encode() for s in sentences]
    end = np.cumsum([len(b) for b in enc]).astype(np.int64)
    begin = np.concatenate([[0], end[:-1]]).astype(np.int64)
    return b"".join(enc), begin, end


def _both(pkg, jg, sentences, emb, first=True):
    text, begin, end = _ranges(sentences)
    d = torch.from_numpy(np.ascontiguousarray(emb)).cuda()
    got = jg.sentence_items(text, begin, end, d.data_ptr(), emb.shape[0], emb.shape[1], first).tobytes()
    want = pkg.abi.json_sentence_items_ranges(text, begin, end, emb, first).tobytes()
    if got != want:
        k = next(i for i, (a, b) in enumerate(zip(got, want)) if a != b) if len(got) == len(want) else min(len(got), len(want))
        raise AssertionError(f"differs at byte {k} of {len(want)} / {len(got)}: {want[max(0, k - 40):k + 40]!r} vs {got[max(0, k - 40):k + 40]!r}")
    return got


def test_items_are_the_host_formatters_bytes(pkg, built):
    rng = np.random.default_rng(0)
    with pkg.abi.JsonGpu(0) as jg:
        # what embeddings look like, at the three hidden sizes, with texts that need every escape
        texts = ["plain", 'quote " backslash \\ slash /', "tab\tnewline\ncr\rbs\bff\f", "ctl \x01\x1f del \x7f", "unicode é 日本 \U0001f600", ""]
        for dim in (384, 768, 128):
            n = 517
            emb = (rng.standard_normal((n, dim)) * rng.choice([1e-3, 0.05, 1.0, 30.0], size=(n, 1))).astype(np.float32)
            doc = _both(pkg, jg, [texts[i % len(texts)] + str(i) for i in range(n)], emb)
            assert doc.startswith(b'{"sentence_text":"plain0","embedding":[') and doc.endswith(b"]}")
        _both(pkg, jg, ["a", "b"], rng.standard_normal((2, 384)).astype(np.float32), first=False)  # continuation chunk: leading comma
        # one sentence, one float
        _both(pkg, jg, ["x"], np.array([[0.5]], np.float32))


def test_special_values_and_layout_edges(pkg, built):
    f = np.float32
    specials = [0.0, -0.0, 1.0, -1.0, 0.1, 0.3, 1e-5, 1e-6, 9.999999e-7, 1e-7, 123456.0, 1e7, 16777216.0, 1e13, 9.999999e12, 1e14, 1e16, 1e21,
                1.5e21, 3.4028235e38, -3.4028235e38, 1.17549435e-38, 1e-45, 1.4e-45, 5e-324, np.inf, -np.inf, np.nan, 0.001234, 12.34,
                1234e7, 0.000001234, 8388608.0, 8388607.5, 0.33333334, 2.0 / 3, 100.0, 1e2, 1e10, 1.0000001, 0.99999994]
    vals = np.array(specials, dtype=np.float32)
    pad = (-len(vals)) % 8
    emb = np.concatenate([vals, np.zeros(pad, f)]).reshape(-1, 8)
    with pkg.abi.JsonGpu(0) as jg:
        _both(pkg, jg, [f"s{i}" for i in range(emb.shape[0])], emb)


def test_bit_pattern_sweep(pkg, built):
    """Every 4099th binary32 bit pattern (1 M values over the whole range: normals, subnormals, infinities, NaNs, both signs)
    and 200 k random patterns: the device digits and layout equal the host's, which is checked exhaustively against
    std::to_chars (scripts/probes/f32_digits_check.cpp)."""
    bits = np.arange(0, 2 ** 32, 4099, dtype=np.uint64).astype(np.uint32)
    rnd = np.random.default_rng(1).integers(0, 2 ** 32, size=200_000, dtype=np.uint64).astype(np.uint32)
    allbits = np.concatenate([bits, rnd])
    allbits = allbits[: len(allbits) // 256 * 256]
    emb = allbits.view(np.float32).reshape(-1, 256)
    with pkg.abi.JsonGpu(0) as jg:
        _both(pkg, jg, ["t"] * emb.shape[0], emb)


def test_items_from_the_embedding_call_on_the_device(pkg, built):
    """forward_device -> items on the device: the text equals formatting the host copy of the same embeddings."""
    from gpu_util import dev_u32
    shape = pkg.weights.TINY
    lengths = pkg.synth.lengths_uniform(300, 2, 64, seed=2)
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 64, seed=3)
    sents = [f"sentence number {i}" for i in range(300)]
    text, begin, end = _ranges(sents)
    with pkg.Embedder.random_init(shape, seed=5) as emb, pkg.abi.JsonGpu(0) as jg:
        out = torch.empty((300, shape.hidden_size), dtype=torch.float32, device="cuda")
        a, b, c = dev_u32(ids), dev_u32(mask), dev_u32(tt)
        emb.forward_device(a.data_ptr(), b.data_ptr(), c.data_ptr(), 300, 64, out.data_ptr())
        torch.cuda.synchronize()
        got = jg.sentence_items(text, begin, end, out.data_ptr(), 300, shape.hidden_size).tobytes()
        want = pkg.abi.json_sentence_items_ranges(text, begin, end, out.cpu().numpy()).tobytes()
    assert got == want
