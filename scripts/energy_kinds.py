"""GPU: board power, SM clock and energy per launch of each large kernel run alone in a loop for a few seconds, with a
cuBLAS bf16 GEMM beside them as the J / TFLOP yardstick. The c2 step is bound by the board's power limit, so energy per
launch - not time alone - is what a kernel costs the step.   usage: python scripts/energy_kinds.py [seconds]"""
import os, subprocess, sys, threading, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_package
pkg = load_package()
from gpu_util import dev_i32, cu_of

SECS = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
B, S, H, I = 4096, 128, 384, 1536
T = B * S

def sample(stop, out):
    while not stop.is_set():
        r = subprocess.run(["nvidia-smi", "--query-gpu=power.draw,clocks.sm", "--format=csv,noheader,nounits", "-i", "0"],
                           capture_output=True, text=True).stdout.strip().split(",")
        try: out.append((float(r[0]), float(r[1])))
        except Exception: pass
        time.sleep(0.05)

def run(name, f, flops):
    for _ in range(3): f()
    torch.cuda.synchronize()
    stop, samples = threading.Event(), []
    th = threading.Thread(target=sample, args=(stop, samples)); th.start()
    n, t0 = 0, time.time()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < SECS:
        for _ in range(20): f()
        n += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    stop.set(); th.join()
    ms = e0.elapsed_time(e1) / n
    s = samples[len(samples) // 3:]  # after the ramp
    p = float(np.mean([a for a, _ in s])); c = float(np.median([b for _, b in s]))
    print(f"{name:28s} {ms:8.4f} ms/launch  {p:6.0f} W  {c:5.0f} MHz  {p * ms / 1e3:7.4f} J/launch  "
          f"{flops / ms / 1e9:7.1f} TFLOP/s  {p * ms / 1e3 / (flops / 1e12):6.3f} J/TFLOP" if flops else
          f"{name:28s} {ms:8.4f} ms/launch  {p:6.0f} W  {c:5.0f} MHz  {p * ms / 1e3:7.4f} J/launch", flush=True)

a = torch.randn((8192, 8192), device="cuda").bfloat16(); b = torch.randn((8192, 8192), device="cuda").bfloat16()
run("cuBLAS bf16 8192^3", lambda: torch.matmul(a, b), 2 * 8192 ** 3)
del a, b
shape = pkg.weights.BertShape(512, H, 1, 12, I, 128, 2)
td = torch.bfloat16
x = torch.randn((T, H), device="cuda").to(td)
lo = torch.zeros((T, H), dtype=torch.uint8, device="cuda")
y = torch.empty((T, H), dtype=td, device="cuda"); y_lo = torch.empty((T, H), dtype=torch.uint8, device="cuda")
d_cu = dev_i32(cu_of(np.full(B, S)))
with pkg.Embedder.random_init(shape, seed=1, operands="bf16") as e:
    run("K23 qkv + attention", lambda: e.op_qkv_attention(0, x.data_ptr(), d_cu.data_ptr(), B, T, S, y.data_ptr()),
        2 * T * 3 * H * H + 4 * H * B * S * S)
    run("attention-output GEMM + LN", lambda: e.op_linear(0, 1, x.data_ptr(), x.data_ptr(), T, 0, y.data_ptr(), d_residual_lo=lo.data_ptr(), d_y_lo=y_lo.data_ptr()),
        2 * T * H * H)
    run("fused FFN", lambda: e.op_ffn(0, x.data_ptr(), lo.data_ptr(), T, y.data_ptr(), y_lo.data_ptr()), 4 * T * I * H)
