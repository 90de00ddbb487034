"""GPU: timing of the fused QKV + attention op (K23) at the c2 shape (also the target of ncu captures)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gpu_util import dev_i32, cu_of
B, S, H = 4096, 128, 384
shape = pkg.weights.BertShape(512, H, 1, 12, 1536, 128, 2)
x = torch.randn((B * S, H), device="cuda").bfloat16()
d_cu = dev_i32(cu_of(np.full(B, S)))
ctx = torch.empty((B * S, H), dtype=torch.bfloat16, device="cuda")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
with pkg.Embedder.random_init(shape, seed=1) as e:
    f = lambda: e.op_qkv_attention(0, x.data_ptr(), d_cu.data_ptr(), B, B * S, S, ctx.data_ptr())
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    print(f"qkv_attention {B} x {S}: {e0.elapsed_time(e1) / reps:.3f} ms per launch")
