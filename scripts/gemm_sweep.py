"""GPU: steady-state rate of the tcgen05 GEMM mainloop for several tile widths / K (CUDA events, 20 launches)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
import pynvml
pynvml.nvmlInit(); hnd = pynvml.nvmlDeviceGetHandleByIndex(0)
def run(m, n, k, bn, act=-1, reps=20):
    a = (torch.randn((m, k), device="cuda") * 0.5).bfloat16()
    b = (torch.randn((n, k), device="cuda") * 0.05).bfloat16()
    bias = torch.zeros((n,), device="cuda")
    c = torch.empty((m, n), dtype=torch.bfloat16, device="cuda")
    f = lambda: pkg.abi.op_gemm_bf16(a.data_ptr(), b.data_ptr(), bias.data_ptr(), m, n, k, act, bn, c.data_ptr())
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    clk = pynvml.nvmlDeviceGetClockInfo(hnd, pynvml.NVML_CLOCK_SM)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    tf = 2.0 * m * n * k / (ms * 1e-3) / 1e12
    print(f"M {m:7d} N {n:5d} K {k:5d} BN {bn:3d} act {act:2d}: {ms:8.3f} ms  {tf:7.1f} TFLOP/s  sm clock {clk} MHz  "
          f"= {100*tf/(148*8192*clk*1e6/1e12):.1f}% of MMA peak at that clock", flush=True)
for ns in ('1', '2'):
    os.environ['B200_EMBED_GEMM_NSPLIT'] = ns
    print('NSPLIT', ns)
    for bn in (256, 192, 128):
        run(65536, 3072 if bn != 192 else 2304, 4096, bn)
    run(524288, 1152, 384, 192)
    run(524288, 1536, 384, 256)
os.environ['B200_EMBED_GEMM_NSPLIT'] = '1'
run(524288, 1152, 384, 192)
run(524288, 1536, 384, 256)
run(524288, 1536, 384, 256, act=0)
run(524288, 1536, 384, 128)
run(2097152, 2304, 768, 256, reps=5)
# cuBLAS for the same shapes (library reference point)
for (m, n, k) in ((65536, 3072, 4096), (524288, 1152, 384), (524288, 1536, 384)):
    a = torch.randn((m, k), device="cuda").bfloat16(); b = torch.randn((n, k), device="cuda").bfloat16()
    for _ in range(3): torch.matmul(a, b.t())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): torch.matmul(a, b.t())
    clk = pynvml.nvmlDeviceGetClockInfo(hnd, pynvml.NVML_CLOCK_SM)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"cuBLAS M {m} N {n} K {k}: {ms:.3f} ms {2.0*m*n*k/(ms*1e-3)/1e12:.1f} TFLOP/s clock {clk}", flush=True)
