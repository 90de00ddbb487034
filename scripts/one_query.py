"""GPU: a few single-query calls (target for an ncu launch list)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
with pkg.Embedder.random_init(shape, seed=1234) as emb:
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [32], 32)
    for _ in range(3): emb.forward(ids, mask, tt)
print("ok")
