"""GPU: a few launches of one GEMM shape through b200_op_gemm_bf16 (target for an ncu capture).
usage: python scripts/gemm_one.py M N K BN ACT [launches]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
m, n, k, bn, act = (int(x) for x in sys.argv[1:6])
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 4
a = (torch.randn((m, k), device="cuda") * 0.5).bfloat16()
b = (torch.randn((n, k), device="cuda") * 0.05).bfloat16()
bias = torch.zeros((n,), device="cuda")
c = torch.empty((m, n), dtype=torch.bfloat16, device="cuda")
for _ in range(reps):
    pkg.abi.op_gemm_bf16(a.data_ptr(), b.data_ptr(), bias.data_ptr(), m, n, k, act, bn, c.data_ptr())
torch.cuda.synchronize()
print("ok")
