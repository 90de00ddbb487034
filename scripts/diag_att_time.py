import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gpu_util import dev_i32, cu_of
shape = pkg.weights.MINILM_L6
B, S, H = 4096, 128, 384
qkv = torch.randn((B * S, 3 * H), device="cuda").bfloat16()
d_cu = dev_i32(cu_of(np.full(B, S)))
ctx = torch.empty((B * S, H), dtype=torch.bfloat16, device="cuda")
with pkg.Embedder.random_init(shape, seed=1) as e:
    for dt in (0, 2):
        for _ in range(3): e.op_attention(qkv.data_ptr(), d_cu.data_ptr(), B, B * S, S, dt, ctx.data_ptr())
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): e.op_attention(qkv.data_ptr(), d_cu.data_ptr(), B, B * S, S, dt, ctx.data_ptr())
        e1.record(); torch.cuda.synchronize()
        print("dbg", os.environ.get("B200_EMBED_ATTN_DBG"), "kernel", "tc" if dt == 0 else "mma.sync", "ms", e0.elapsed_time(e1) / 10)
