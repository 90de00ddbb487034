"""GPU: does running the layer stack over slabs that fit the 126 MB L2 beat one pass over the whole batch?
cfg.max_tokens_per_launch splits a call into sentence chunks that each go through all layers (slab-major order), so the
activations a slab writes (ctx, a, x: 10 H bytes per token) are re-read from L2 instead of HBM. c2 (4096 x 128), MiniLM-L6.
usage: python scripts/slab_ab.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
B, S = 4096, 128
ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, pkg.synth.lengths_dense(B, S), S, seed=42)
as_dev = lambda a: torch.from_numpy(a.view(np.int32)).cuda()
d = [as_dev(a) for a in (ids, mask, tt)]
out = torch.empty((B, shape.hidden_size), dtype=torch.float32, device="cuda")
st = torch.cuda.Stream()
res = {}
for rounds in range(3):
    for tiles in (148, 296, 592, 1184, 4096):
        with pkg.Embedder.random_init(shape, seed=1, operands="bf16", max_tokens_per_launch=tiles * 128) as e:
            f = lambda: e.forward_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), B, S, out.data_ptr(), stream=st.cuda_stream)
            with torch.cuda.stream(st):
                for _ in range(3): f()
                st.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(10): f()
                e1.record(st); st.synchronize()
            res.setdefault(tiles, []).append(e0.elapsed_time(e1) / 10)
for tiles, v in res.items():
    print(f"slab of {tiles:5d} tiles ({tiles * 128 * 3840 / 1e6:7.1f} MB of activations): ms per 4096 sentences " + " ".join(f"{x:.3f}" for x in v))
