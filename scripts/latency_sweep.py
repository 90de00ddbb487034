"""GPU: b200_embed_forward latency vs batch size for the current schedule (run under different env switches)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
out = []
with pkg.Embedder.random_init(shape, seed=1234) as emb:
    for batch in (1, 8, 32, 64, 128, 256, 512, 1024):
        seq = 128
        ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [seq] * batch, seq)
        for _ in range(10): emb.forward(ids, mask, tt)
        ts = []
        for _ in range(50):
            t0 = time.perf_counter(); emb.forward(ids, mask, tt); ts.append(time.perf_counter() - t0)
        out.append(f"{batch}:{np.median(ts) * 1e3:.3f}")
print(" ".join(out))
