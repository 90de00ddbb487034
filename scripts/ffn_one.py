"""GPU: launches of the fused FFN op at the c2 shape (timing, or the target of an ncu capture)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
shape = pkg.weights.BertShape(512, 384, 1, 12, 1536, 64, 2)
with pkg.Embedder.random_init(shape, seed=4) as emb:
    x = torch.randn((rows, 384), device="cuda").bfloat16(); xl = torch.zeros_like(x)
    y = torch.empty_like(x); yl = torch.empty_like(x)
    f = lambda: emb.op_ffn(0, x.data_ptr(), xl.data_ptr(), rows, y.data_ptr(), yl.data_ptr())
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    print(f"fused ffn rows {rows}: {e0.elapsed_time(e1) / reps:.3f} ms per launch")
