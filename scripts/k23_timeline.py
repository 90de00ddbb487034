"""GPU: run K23 once on the c2 shape with a -DQA_TIMELINE build (python codename-symbiont_b200/build.py --suffix _tl
--flags=-DQA_TIMELINE) and let the kernel print the cycles one thread per role of block 0 spent in each phase.
usage: B200_EMBED_LIB=codename-symbiont_b200/libb200_embed_tl.so python scripts/k23_timeline.py [batch] [seq]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_package
pkg = load_package()
from gpu_util import dev_i32, cu_of

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
H = 384
shape = pkg.weights.BertShape(512, H, 1, 12, 1536, 128, 2)
x = torch.randn((B * S, H), device="cuda").to(torch.bfloat16)
d_cu = dev_i32(cu_of(np.full(B, S)))
y = torch.empty((B * S, H), dtype=torch.bfloat16, device="cuda")
with pkg.Embedder.random_init(shape, seed=1, operands="bf16") as e:
    for i in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        e.op_qkv_attention(0, x.data_ptr(), d_cu.data_ptr(), B, B * S, S, y.data_ptr())
        e1.record(); torch.cuda.synchronize()
        print(f"launch {i}: {e0.elapsed_time(e1):.4f} ms", flush=True)
