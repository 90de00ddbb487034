"""CPU: the oracle against the committed golden vectors (HuggingFace fp64 on identical weights),
plus the size-independent properties the GPU tests rely on. Reference: embedding_generator.rs:134-223."""
import numpy as np
import pytest

from conftest import load_golden, cos_rows
from oracle.oracle import Oracle


@pytest.fixture(scope="module")
def tiny(pkg, built):
    shape = pkg.weights.TINY
    w = pkg.weights.init_weights(shape, 1234)
    return shape, w, Oracle(shape, w, "f32"), Oracle(shape, w, "f64")


@pytest.mark.parametrize("name", ["tiny_dense", "tiny_ragged", "tiny_single"])
def test_oracle_matches_hf_golden(pkg, built, name):
    g = load_golden(name)
    shape = pkg.weights.SHAPES[str(g["shape_key"])]
    w = pkg.weights.init_weights(shape, int(g["seed"]))
    for prec, tol in (("f32", 2e-6), ("f64", 2e-7)):
        got = Oracle(shape, w, prec).generate(g["ids"], g["type_ids"], g["mask"], chunk=8)
        assert got.shape == g["expected_hf_f64"].shape
        assert cos_rows(got, g["expected_hf_f64"]).min() >= 0.999999
        assert np.abs(got - g["expected_hf_f64"]).max() <= tol


def test_oracle_minilm_golden(pkg, built):
    g = load_golden("minilm_b8_s128")
    shape = pkg.weights.MINILM_L6
    w = pkg.weights.init_weights(shape, int(g["seed"]))
    got = Oracle(shape, w, "f32").generate(g["ids"], g["type_ids"], g["mask"], chunk=8)
    assert cos_rows(got, g["expected_hf_f64"]).min() >= 0.999999
    assert np.abs(got - g["expected_hf_f64"]).max() <= 5e-6


def test_empty_in_empty_out(tiny):
    shape, w, o32, _ = tiny
    z = np.zeros((0, 16), dtype=np.uint32)
    assert o32.generate(z, z, z).shape == (0, shape.hidden_size)  # embedding_generator.rs:135-137


def test_padding_and_chunking_do_not_change_results(pkg, tiny):
    """Padded positions never reach the output: the licence for the varlen CUDA path."""
    shape, w, o32, _ = tiny
    lengths = [3, 9, 16, 2, 12, 7, 16, 5, 11]
    ids16, m16, t16 = pkg.synth.padded_batch(shape.vocab_size, lengths, 16, seed=3)
    ids64 = np.zeros((len(lengths), 64), np.uint32); ids64[:, :16] = ids16
    m64 = np.zeros_like(ids64); m64[:, :16] = m16
    ids64[m64 == 0] = 77  # the pad id is irrelevant too
    a = o32.generate(ids16, t16, m16, chunk=8)
    b = o32.generate(ids64, np.zeros_like(ids64), m64, chunk=8)
    c = o32.generate(ids16, t16, m16, chunk=3)
    assert np.abs(a - b).max() <= 1e-6
    assert np.array_equal(a, c)


def test_pool_formula_and_l2(tiny):
    shape, w, o32, _ = tiny
    rng = np.random.default_rng(0)
    hid = rng.standard_normal((3, 8, shape.hidden_size)).astype(np.float32)
    mask = np.array([[1] * 8, [1] * 3 + [0] * 5, [1] + [0] * 7], np.uint32)
    got = o32.mean_pool(hid, mask)
    want = (hid * mask[:, :, None]).sum(1) / (mask.sum(1, keepdims=True) + np.float32(1e-9))
    assert np.abs(got - want).max() <= 1e-6
    n = o32.l2_normalize(got)
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-6)
    assert cos_rows(n, got).min() >= 0.9999999


def test_out_of_range_inputs_rejected(tiny):
    shape, w, o32, _ = tiny
    ids = np.full((1, 4), shape.vocab_size, np.uint32)
    one = np.ones((1, 4), np.uint32)
    with pytest.raises(ValueError):
        o32.generate(ids, np.zeros_like(ids), one)
    with pytest.raises(ValueError):
        o32.generate(np.ones((1, 65), np.uint32), np.zeros((1, 65), np.uint32), np.ones((1, 65), np.uint32))
