"""Throughput of the JSON wire-format stage (f3). Usage: python scripts/json_bench.py [n_sentences] [dim]"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
abi = importlib.import_module("codename-symbiont_b200.abi")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 384
rng = np.random.default_rng(0)
emb = (rng.standard_normal((n, dim)) * 0.05).astype(np.float32)
text = b"The quick brown fox jumps over the lazy dog again." * n
begin = np.arange(n, dtype=np.int64) * 50
end = begin + 50
for threads in (1, 0):
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        out = abi.json_text_with_embeddings("doc", "http://x", (text, begin, end), emb, "m", 1, threads=threads, as_array=True)
        best = min(best, time.perf_counter() - t0)
    print(f"threads {threads or os.cpu_count()}: {best * 1e3:.1f} ms, {n / best:.0f} sentences/s, {len(out) / best / 1e6:.0f} MB/s, {best / (n * dim) * 1e9 * (threads or os.cpu_count()):.1f} ns per float per core")
