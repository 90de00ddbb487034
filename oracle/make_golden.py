"""Pins the oracle (and, through it, the CUDA path) in THIS repo -- run in the build container.

The reference pins nothing for the embedding path and cannot be
compiled here (Rust, un-vendored candle crates). candle's `BertModel` is a port of HuggingFace
BERT, so the independent check is `transformers.BertModel` (eager attention, no pooler, CPU
fp32 and fp64) on the identical weights and token ids, followed by the reference's pooling
formula (embedding_generator.rs:201-207) written in numpy.

  python oracle/make_golden.py            # checks oracle == HF, then (re)writes tests/golden/*.npz

Each fixture stores the shape, the weight seed (weights are regenerated from the spec in
codename-symbiont_b200/weights.py), the padded token ids / mask / type ids, and the expected
pooled [N,H] f32 embeddings from HF fp64 and from the f32 oracle.
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") != os.path.join(ROOT, "oracle")]
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
from oracle.oracle import Oracle  # noqa: E402


def hf_forward(shape, weights, ids, mask, type_ids, dtype):
    import torch
    from transformers import BertConfig, BertModel

    act = {0: "gelu", 1: "gelu_new", 2: "relu"}[shape.hidden_act]
    cfg = BertConfig(vocab_size=shape.vocab_size, hidden_size=shape.hidden_size,
                     num_hidden_layers=shape.num_hidden_layers, num_attention_heads=shape.num_attention_heads,
                     intermediate_size=shape.intermediate_size, hidden_act=act,
                     max_position_embeddings=shape.max_position_embeddings, type_vocab_size=shape.type_vocab_size,
                     layer_norm_eps=shape.layer_norm_eps, hidden_dropout_prob=0.0,
                     attention_probs_dropout_prob=0.0, attn_implementation="eager")
    model = BertModel(cfg, add_pooling_layer=False).eval()
    sd = {k: torch.from_numpy(v.copy()) for k, v in weights.items()}
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("position_ids" in m for m in missing), missing
    model = model.to(dtype)
    with torch.no_grad():
        S = ids.shape[1]
        out = model(input_ids=torch.from_numpy(ids.astype(np.int64)),
                    attention_mask=torch.from_numpy(mask.astype(np.int64)),
                    token_type_ids=torch.from_numpy(type_ids.astype(np.int64)),
                    position_ids=torch.arange(S)[None, :]).last_hidden_state
    return out.numpy()


def pool_numpy(hidden, mask):
    m = mask.astype(hidden.dtype)[:, :, None]
    return ((hidden * m).sum(1) / (m.sum(1) + np.float32(1e-9))).astype(np.float32)


def cos_rows(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return (a * b).sum(1) / (np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1))


CASES = {
    # name: (shape key, seed, lengths, pad_to)
    "tiny_dense": ("tiny", 1234, [16] * 6, 16),
    "tiny_ragged": ("tiny", 1234, [2, 3, 7, 8, 9, 31, 32, 33, 63, 64, 17, 5], 64),
    "tiny_single": ("tiny", 99, [11], 64),
}


def main():
    outdir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(outdir, exist_ok=True)
    for name, (skey, seed, lengths, pad_to) in CASES.items():
        shape = pkg.weights.SHAPES[skey]
        w = pkg.weights.init_weights(shape, seed)
        ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, pad_to, seed=42)
        if shape.type_vocab_size > 1:  # exercise the token-type table too
            tt = ((np.arange(pad_to)[None, :] >= (np.asarray(lengths)[:, None] // 2)) & (mask > 0)).astype(np.uint32)
        h64 = hf_forward(shape, w, ids, mask, tt, __import__("torch").float64)
        h32 = hf_forward(shape, w, ids, mask, tt, __import__("torch").float32)
        hf64, hf32 = pool_numpy(h64, mask), pool_numpy(h32, mask)
        o32 = Oracle(shape, w, "f32").generate(ids, tt, mask, chunk=8)
        o64 = Oracle(shape, w, "f64").generate(ids, tt, mask, chunk=8)
        for label, a in (("oracle_f32", o32), ("oracle_f64", o64), ("hf_f32", hf32)):
            c = cos_rows(a, hf64).min()
            e = np.abs(a - hf64).max()
            print(f"{name:12s} {label:10s} vs hf_f64: min cos {c:.9f}  max abs {e:.3e}")
            assert c >= 0.999999 and e <= 1e-5, "oracle and HuggingFace disagree"
        np.savez_compressed(os.path.join(outdir, name + ".npz"), shape_key=skey, seed=seed, ids=ids, mask=mask,
                            type_ids=tt, expected_hf_f64=hf64, expected_oracle_f32=o32)
    # one mid-size check on the BASELINE config-1 shape (not stored: weights are 90 MB)
    shape = pkg.weights.MINILM_L6
    w = pkg.weights.init_weights(shape, 1234)
    lengths = pkg.synth.lengths_uniform(8, 8, 128, seed=7)
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 128, seed=42)
    hf64 = pool_numpy(hf_forward(shape, w, ids, mask, tt, __import__("torch").float64), mask)
    o32 = Oracle(shape, w, "f32").generate(ids, tt, mask, chunk=8)
    c, e = cos_rows(o32, hf64).min(), np.abs(o32 - hf64).max()
    print(f"minilm-l6 B8xS128 oracle_f32 vs hf_f64: min cos {c:.9f}  max abs {e:.3e}")
    assert c >= 0.999999 and e <= 1e-5
    np.savez_compressed(os.path.join(outdir, "minilm_b8_s128.npz"), shape_key="minilm-l6", seed=1234, ids=ids,
                        mask=mask, type_ids=tt, expected_hf_f64=hf64, expected_oracle_f32=o32)


if __name__ == "__main__":
    main()
