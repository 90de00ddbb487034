"""GPU: latency of the query path: one short sentence through b200_embed_forward with host buffers."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
with pkg.Embedder.random_init(shape, seed=1234) as emb:
    for batch, seq in ((1, 16), (1, 32), (1, 128), (8, 32), (8, 512)):
        ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [seq] * batch, seq)
        for _ in range(20): emb.forward(ids, mask, tt)
        ts = []
        for _ in range(200):
            t0 = time.perf_counter(); emb.forward(ids, mask, tt); ts.append(time.perf_counter() - t0)
        ts = np.array(ts) * 1e3
        print(f"batch {batch} x seq {seq}: median {np.median(ts):.3f} ms  p10 {np.percentile(ts, 10):.3f}  p90 {np.percentile(ts, 90):.3f}", flush=True)
