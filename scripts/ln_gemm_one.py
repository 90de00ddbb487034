"""GPU: timing of the LayerNorm-epilogue GEMMs (attn_out: which=1, FFN2: which=3) through b200_op_linear.
usage: python scripts/ln_gemm_one.py [rows] [reps]   (B200_EMBED_GEMM_DBG is read once per process)"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
H = int(os.environ.get("LN_H", "384")); shape = pkg.weights.BertShape(512, H, 1, 12, 4 * H, 64, 2)
with pkg.Embedder.random_init(shape, seed=4) as emb:
    for which, k in ((1, H), (3, 4 * H)):
        x = torch.randn((rows, k), device="cuda").bfloat16()
        r = torch.randn((rows, H), device="cuda").bfloat16(); rl = torch.zeros_like(r)
        y = torch.empty_like(r); yl = torch.empty_like(r)
        f = lambda: emb.op_linear(0, which, x.data_ptr(), r.data_ptr(), rows, 0, y.data_ptr(), rl.data_ptr(), yl.data_ptr())
        f(); f(); torch.cuda.synchronize()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps): f()
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / reps)
        print(f"dbg={os.environ.get('B200_EMBED_GEMM_DBG', '0')} which={which} K={k}: {best:.3f} ms", flush=True)
