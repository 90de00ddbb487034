"""GPU: randomised end-to-end parity sweep. Every trial draws a batch (size, padded length, length distribution), runs it
through (a) the fused schedule forced on (K23 + K2f), (b) the unfused schedule, (c) whatever the engine picks by itself,
and - for small trials - the f32 C oracle; reports the worst differences. Usage: python scripts/fuzz_parity.py [trials] [seed] [f16|bf16]"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from __graft_entry__ import load_package
pkg = load_package()
from oracle.oracle import Oracle

trials = int(sys.argv[1]) if len(sys.argv) > 1 else 150
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
operands = sys.argv[3] if len(sys.argv) > 3 else "f16"
shape = pkg.weights.MINILM_L6
rng = np.random.default_rng(seed)

def make(env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return pkg.Embedder.random_init(shape, seed=77, operands=operands)
    finally:
        for k, v in old.items():
            os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)

fused = make({"B200_EMBED_FFN_FUSED": "1", "B200_EMBED_QKV_ATTN": "1"})
plain = make({"B200_EMBED_FFN_FUSED": "0", "B200_EMBED_QKV_ATTN": "0"})
auto = make({})
w = plain.export_weights()
oracle = Oracle(shape, w, "f32")
def cos(a, b):
    return (a * b).sum(1) / np.maximum(np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1), 1e-30)
worst = dict(fp=0.0, fa=0.0, cos=1.0, orc=0.0, orc_cos=1.0)
t0 = time.time()
for t in range(trials):
    seq = int(rng.choice([2, 3, 8, 16, 31, 32, 33, 64, 65, 100, 127, 128]))
    batch = int(rng.choice([1, 2, 3, 7, 31, 32, 33, 64, 100, 255, 256, 257, 300, 511, 700, 1500]))
    kind = rng.integers(0, 5)
    if kind == 0: lengths = np.full(batch, seq)
    elif kind == 1: lengths = rng.integers(2, seq + 1, batch)
    elif kind == 2: lengths = np.where(rng.random(batch) < 0.8, 2, seq)
    elif kind == 3: lengths = np.minimum(pkg.synth.lengths_corpus(batch, seed=int(rng.integers(1 << 30))), seq).clip(2, seq)
    else: lengths = np.sort(rng.integers(2, seq + 1, batch))[::-1].copy()
    lengths = np.maximum(lengths, 2)
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, seq, seed=int(rng.integers(1 << 30)))
    a, b, c = fused.forward(ids, mask, tt), plain.forward(ids, mask, tt), auto.forward(ids, mask, tt)
    assert np.isfinite(a).all() and np.isfinite(b).all() and np.isfinite(c).all(), (t, batch, seq, kind)
    d_fp, d_fa = float(np.abs(a - b).max()), float(min(np.abs(c - a).max(), np.abs(c - b).max()))
    cmin = float(min(cos(a, b).min(), cos(c, b).min()))
    worst["fp"] = max(worst["fp"], d_fp); worst["fa"] = max(worst["fa"], d_fa); worst["cos"] = min(worst["cos"], cmin)
    if d_fp > 6e-3 or cmin < 0.9999:
        print(f"trial {t}: batch {batch} seq {seq} kind {kind}: fused-plain {d_fp:.2e} cos {cmin:.6f}  <-- LARGE", flush=True)
    if batch * seq <= 4096 and t % 3 == 0:
        want = oracle.generate(ids, tt, mask, chunk=8)
        for x in (a, b):
            worst["orc"] = max(worst["orc"], float(np.abs(x - want).max())); worst["orc_cos"] = min(worst["orc_cos"], float(cos(x, want).min()))
print(f"operands {operands}: {trials} trials in {time.time() - t0:.0f}s: max |fused - unfused| {worst['fp']:.2e}, auto equals one of them within {worst['fa']:.1e}, "
      f"min cosine between schedules {worst['cos']:.7f}; vs f32 oracle: max abs {worst['orc']:.2e}, min cosine {worst['orc_cos']:.7f}")
