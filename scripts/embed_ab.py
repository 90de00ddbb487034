"""GPU: time K1 (embeddings + LayerNorm) and K5 (pooling) alone on the c2 shape for several library builds.
usage: python scripts/embed_ab.py lib_a.so lib_b.so ..."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import os, sys
import numpy as np, torch
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
from __graft_entry__ import load_package
pkg = load_package()
from gpu_util import dev_i32, dev_u32, cu_of
B, S = 4096, 128
shape = pkg.weights.MINILM_L6
H = shape.hidden_size
ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, pkg.synth.lengths_dense(B, S), S, seed=42)
pi, pt, cu = pkg.synth.pack(ids, mask, tt)
d_pi, d_pt, d_cu = dev_u32(pi), dev_u32(pt), dev_i32(cu)
out = torch.empty((B * S, H), dtype=torch.bfloat16, device="cuda"); out_lo = torch.empty((B * S, H), dtype=torch.uint8, device="cuda")
pooled = torch.empty((B, H), dtype=torch.float32, device="cuda")
with pkg.Embedder.random_init(shape, seed=1, operands="bf16") as e:
    fs = {"embed": lambda: e.op_embed_layernorm(d_pi.data_ptr(), d_pt.data_ptr(), d_cu.data_ptr(), B, B * S, 0, out.data_ptr(), out_lo.data_ptr()),
          "pool": lambda: e.op_mean_pool(out.data_ptr(), d_cu.data_ptr(), B, 0, 0, pooled.data_ptr(), d_hidden_lo=out_lo.data_ptr())}
    res = {}
    for name, f in fs.items():
        for _ in range(3): f()
        torch.cuda.synchronize()
        ts = []
        for _ in range(30):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); f(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort(); res[name] = ts[len(ts) // 2]
    print(" ".join(f"{k} {v:.4f} ms" for k, v in res.items()))
"""
for r in range(2):
    for l in sys.argv[1:]:
        env = dict(os.environ, B200_EMBED_LIB=os.path.abspath(l))
        out = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}], env=env, capture_output=True, text=True)
        print(os.path.basename(l), out.stdout.strip().splitlines()[-1] if out.stdout.strip() else "FAILED " + out.stderr[-400:])
