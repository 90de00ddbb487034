"""GPU: K=384 GEMM timing under the experiment switches (B200_EMBED_GEMM_PAIR / _WIDE_STORE / _DBG).
Configurations are interleaved round-robin (power / clock state drifts during a run) and the best round is reported."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
M = 524288
shapes = [(M, 1152, 384, 192, -1), (M, 1536, 384, 256, -1), (M, 1536, 384, 256, 0)]
configs = [dict(PAIR=p, DBG=d, WIDE_STORE=w) for p in ("0",) for d, w in (("0", "0"), ("0", "1"), ("0", "2"), ("8", "0"), ("256", "0"))]
for (m, n, k, bn, act) in shapes:
    a = (torch.randn((m, k), device="cuda") * 0.5).bfloat16()
    b = (torch.randn((n, k), device="cuda") * 0.05).bfloat16()
    bias = torch.randn((n,), device="cuda")
    c = torch.empty((m, n), dtype=torch.bfloat16, device="cuda")
    f = lambda: pkg.abi.op_gemm_bf16(a.data_ptr(), b.data_ptr(), bias.data_ptr(), m, n, k, act, bn, c.data_ptr())
    rows = torch.cat([torch.arange(0, 300), torch.arange(m - 300, m)]).cuda()
    ref = a[rows].float() @ b.float().t() + bias
    if act == 0: ref = torch.nn.functional.gelu(ref)
    best = [1e9] * len(configs); err = [0.0] * len(configs)
    for rnd in range(4):
        for ci, cfg in enumerate(configs):
            for kk, vv in cfg.items(): os.environ["B200_EMBED_GEMM_" + kk] = vv
            if rnd == 0:
                c.zero_(); f(); torch.cuda.synchronize()
                if cfg["DBG"] == "0": err[ci] = (c[rows].float() - ref).abs().max().item()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10): f()
            e1.record(); torch.cuda.synchronize()
            best[ci] = min(best[ci], e0.elapsed_time(e1) / 10)
    for ci, cfg in enumerate(configs):
        tf = 2.0 * m * n * k / (best[ci] * 1e-3) / 1e12
        print(f"M {m} N {n:5d} BN {bn:3d} act {act:2d} pair={cfg['PAIR']} dbg={cfg['DBG']} wide={cfg['WIDE_STORE']}: {best[ci]:7.3f} ms {tf:7.1f} TF/s  err {err[ci]:.4f}", flush=True)
