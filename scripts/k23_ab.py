"""GPU: same-box A/B timing of one kernel across library builds. Each variant (a libb200_embed*.so, built with
`python codename-symbiont_b200/build.py --suffix _x --flags "-D..."`) is timed in its own subprocess, variants interleaved
over several rounds so that clock / power drift hits all of them alike; prints min and median ms per launch.
usage: python scripts/k23_ab.py [--op qkv|ffn|ao|step] [--rounds 4] lib_a.so lib_b.so ..."""
import argparse, os, statistics, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import os, sys
import numpy as np, torch
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
from __graft_entry__ import load_package
pkg = load_package()
from gpu_util import dev_i32, cu_of
op, reps = sys.argv[1], int(sys.argv[2])
B, S, H = 4096, 128, 384
td = torch.bfloat16
if op == "step":
    shape = pkg.weights.MINILM_L6
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, pkg.synth.lengths_dense(B, S), S, seed=42)
    as_dev = lambda a: torch.from_numpy(a.view(np.int32)).cuda()
    d = [as_dev(a) for a in (ids, mask, tt)]
    out = torch.empty((B, H), dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream()
    with pkg.Embedder.random_init(shape, seed=1, operands="bf16") as e:
        f = lambda: e.forward_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), B, S, out.data_ptr(), stream=st.cuda_stream)
        with torch.cuda.stream(st):
            for _ in range(2): f()
            st.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(reps): f()
            e1.record(st); st.synchronize()
        print(e0.elapsed_time(e1) / reps)
else:
    shape = pkg.weights.BertShape(512, H, 1, 12, 1536, 128, 2)
    x = torch.randn((B * S, H), device="cuda").to(td)
    lo = torch.zeros_like(x)
    d_cu = dev_i32(cu_of(np.full(B, S)))
    y = torch.empty((B * S, H), dtype=td, device="cuda"); y_lo = torch.empty_like(y)
    with pkg.Embedder.random_init(shape, seed=1, operands="bf16") as e:
        if op == "qkv":
            f = lambda: e.op_qkv_attention(0, x.data_ptr(), d_cu.data_ptr(), B, B * S, S, y.data_ptr())
        elif op == "ffn":
            f = lambda: e.op_ffn(0, x.data_ptr(), lo.data_ptr(), B * S, y.data_ptr(), y_lo.data_ptr())
        else:
            f = lambda: e.op_linear(0, 1, x.data_ptr(), x.data_ptr(), B * S, 0, y.data_ptr(), d_residual_lo=lo.data_ptr(), d_y_lo=y_lo.data_ptr())
        for _ in range(3): f()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); f(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        print(ts[len(ts) // 2])
"""

ap = argparse.ArgumentParser()
ap.add_argument("--op", default="qkv")
ap.add_argument("--rounds", type=int, default=4)
ap.add_argument("--reps", type=int, default=30)
ap.add_argument("libs", nargs="+")
a = ap.parse_args()
res = {l: [] for l in a.libs}
for r in range(a.rounds):
    for l in a.libs:
        env = dict(os.environ, B200_EMBED_LIB=os.path.abspath(l))
        out = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}, a.op, str(a.reps)], env=env, capture_output=True, text=True)
        try:
            res[l].append(float(out.stdout.strip().splitlines()[-1]))
        except Exception:
            print("FAILED", l, out.stderr[-500:])
for l, v in res.items():
    if v:
        print(f"{a.op:5s} {os.path.basename(l):40s} min {min(v):.4f} median {statistics.median(v):.4f} ms  ({' '.join(f'{x:.3f}' for x in v)})")
