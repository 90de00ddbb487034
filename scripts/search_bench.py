"""f4: device time of the cosine scan (b200_index_search) against the HBM roofline: 4 * dim bytes per vector and pass.
Usage: python scripts/search_bench.py [n_vectors] [dim]"""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
abi = importlib.import_module("codename-symbiont_b200.abi")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 384
peak = 6444.1
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
g = torch.Generator(device="cuda").manual_seed(0)
ix = abi.Index(dim, capacity_hint=n)
step = 250_000
for i in range(0, n, step):
    m = min(step, n - i)
    v = torch.randn((m, dim), device="cuda", generator=g)
    ix.add_device(v.data_ptr(), m)
rng = np.random.default_rng(0)
for stage in ("one-stage", "two-stage"):
  if stage == "two-stage":
      ix.enable_prescan()
      print("with the bf16 copy (2 * dim bytes per vector and pass, candidates rescored from f32):")
  for q, k in ((1, 10), (4, 10), (1, 100), (16, 10)):
      qs = rng.standard_normal((q, dim)).astype(np.float32)
      ix.search(qs, k)
      scan, wall = [], []
      for _ in range(5):
          t0 = time.perf_counter(); ix.search(qs, k); wall.append(time.perf_counter() - t0); scan.append(ix.last_scan_ms)
      passes = (q + 3) // 4
      sm = float(np.median(scan)); wm = float(np.median(wall)) * 1e3
      gbs = passes * n * dim * (2 if (stage == 'two-stage' and q == 1 and k <= 64) else 4) / (sm * 1e-3) / 1e9  # the copy serves single-query passes
      print(f"n {n} dim {dim} q {q} k {k}: scan {sm:.3f} ms = {gbs:.0f} GB/s ({gbs / peak:.2f} of {peak:.0f}), call {wm:.3f} ms, {q / (wm * 1e-3):.0f} queries/s")
