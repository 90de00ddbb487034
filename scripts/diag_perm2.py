import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
lengths = pkg.synth.lengths_uniform(1024, 8, 128, seed=3)
ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 128)
perm = np.random.default_rng(0).permutation(len(lengths))
with pkg.Embedder.random_init(shape, seed=1234) as e:
    base = e.forward(ids, mask, tt)
    c = e.forward(ids[perm], mask[perm], tt[perm])
    d = e.forward(ids[perm], mask[perm], tt[perm])
    a2 = e.forward(ids, mask, tt)
    print("perm==base[perm]", np.array_equal(c, base[perm]), "perm twice equal", np.array_equal(c, d), "base twice equal", np.array_equal(base, a2),
          "maxdiff c-base[perm]", np.abs(c - base[perm]).max(), "rows", int((np.abs(c - base[perm]).max(1) > 0).sum()))
    diff_rows = np.nonzero(np.abs(c - base[perm]).max(1) > 0)[0]
    print("differing rows (in permuted order):", diff_rows[:30], "their lengths", lengths[perm][diff_rows[:30]])
    print("orig index of differing:", perm[diff_rows[:30]])
