"""GPU: where the K=384 GEMM time goes. Debug bits (B200_EMBED_GEMM_DBG): 1 no TMA stores, 8 mainloop only,
16 prefetch A two tiles ahead into L2, 32 evict-first stores, 64 evict-last W loads, 128 every CTA prefetches."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
def run(m, n, k, bn, act=-1, reps=20, tag=""):
    a = (torch.randn((m, k), device="cuda") * 0.5).bfloat16()
    b = (torch.randn((n, k), device="cuda") * 0.05).bfloat16()
    bias = torch.zeros((n,), device="cuda")
    c = torch.empty((m, n), dtype=torch.bfloat16, device="cuda")
    f = lambda: pkg.abi.op_gemm_bf16(a.data_ptr(), b.data_ptr(), bias.data_ptr(), m, n, k, act, bn, c.data_ptr())
    for _ in range(5): f()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): f()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    tf = 2.0 * m * n * k / (best * 1e-3) / 1e12
    print(f"{tag:22s} N {n:5d} K {k:4d} BN {bn:3d} act {act:2d}: {best:7.3f} ms {tf:7.1f} TF/s", flush=True)
M = 524288
run(M, 1536, 384, 256, tag="warm")
for dbg in (0, 16, 16 + 128, 32, 64, 16 + 32, 16 + 32 + 64, 8, 8 + 16, 1, 1 + 16):
    for wide in ("0", "1"):
        if wide == "1" and (dbg & 9): continue
        os.environ["B200_EMBED_GEMM_DBG"] = str(dbg)
        os.environ["B200_EMBED_GEMM_WIDE_STORE"] = wide
        tag = f"dbg={dbg} wide={wide}"
        run(M, 1152, 384, 192, tag=tag)
        run(M, 1536, 384, 256, tag=tag)
        run(M, 1536, 384, 256, act=0, tag=tag)
