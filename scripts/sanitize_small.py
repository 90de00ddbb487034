"""GPU: one small pass over every kernel of the path, meant to run under compute-sanitizer
(scripts/gpu_sanitize.sh: memcheck, racecheck, synccheck; summaries go to profiles/). Shapes are tiny on purpose."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
os.environ["B200_EMBED_GRAPHS"] = "0"
for fused in ("0", "1"):
    os.environ["B200_EMBED_FFN_FUSED"] = os.environ["B200_EMBED_QKV_ATTN"] = fused
    for operands in ("f16", "bf16"):
        for shape, lengths, S in ((pkg.weights.BertShape(512, 384, 1, 12, 1536, 128, 2), [2, 7, 33, 64, 65, 100, 128, 128, 5, 3], 128),
                                  (pkg.weights.BertShape(512, 768, 1, 12, 3072, 256, 2), [2, 200, 129, 64], 256),
                                  (pkg.weights.TINY, [2, 3, 17, 64], 64)):
            ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, S, seed=1)
            with pkg.Embedder.random_init(shape, seed=3, operands=operands) as emb:
                out = emb.forward(ids, mask, tt)
                assert np.isfinite(out).all()
                if fused == "0" and operands == "f16":
                    out32 = emb.forward(ids, mask, tt, flags=pkg.FLAG_FP32 | pkg.FLAG_NORMALIZE)
                    assert np.isfinite(out32).all()
            print("ok", fused, operands, shape.hidden_size, flush=True)
print("sanitize_small: all passes finished")
