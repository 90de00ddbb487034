"""GPU: per-kernel device time of one small call (the library's own event profile), to see what a single query spends."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
with pkg.Embedder.random_init(shape, seed=1234) as emb:
    for batch, seq, real in ((1, 128, 20), (1, 128, 128), (8, 128, 128)):
        ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [real] * batch, seq)
        for _ in range(5): emb.forward(ids, mask, tt)
        emb.profile(True); emb.profile_read(reset=True)
        for _ in range(20): emb.forward(ids, mask, tt)
        st = emb.profile_read(reset=True); emb.profile(False)
        tot = sum(v["total_ms"] for v in st.values()) / 20
        print(f"batch {batch} seq {seq} real {real}: sum of kernels {tot*1e3:.0f} us:", {k: f'{v["total_ms"]/v["launches"]*1e3:.1f}us x{v["launches"]//20}' for k, v in st.items() if v["launches"]})
