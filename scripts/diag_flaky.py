import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
lengths = pkg.synth.lengths_uniform(1024, 8, 128, seed=3)
ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 128)
perm = np.random.default_rng(0).permutation(len(lengths))
wide = lambda a: np.concatenate([a, np.zeros((a.shape[0], 64), np.uint32)], axis=1)
with pkg.Embedder.random_init(shape, seed=1234) as e:
    base = e.forward(ids, mask, tt)
    for rep in range(6):
        for name, args, ref in (("base", (ids, mask, tt), base), ("perm", (ids[perm], mask[perm], tt[perm]), base[perm]),
                                ("wide", (wide(ids), wide(mask), wide(tt)), base)):
            out = e.forward(*args)
            d = np.abs(out - ref).max(1)
            nz = np.nonzero(d > 0)[0]
            if len(nz):
                print(f"rep {rep} {name}: {len(nz)} rows differ, max {d.max():.3e}; rows {nz[:12]} lens {np.asarray(lengths if name != 'perm' else lengths[perm])[nz[:12]]}")
    print("done", os.environ.get("B200_EMBED_ATTENTION"))
