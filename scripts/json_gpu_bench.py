"""GPU: rate of the device JSON formatter (b200_jsongpu_sentence_items) against the host one, 384-d embeddings, output in pinned
memory, D2H of the text included. usage: python scripts/json_gpu_bench.py [sentences]"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
dim = 384
rng = np.random.default_rng(0)
emb = (rng.standard_normal((n, dim)) * 0.05).astype(np.float32)
sents = [("sentence %d " % i) * 6 for i in range(n)]
enc = [s.encode() for s in sents]
end = np.cumsum([len(b) for b in enc]).astype(np.int64); begin = np.concatenate([[0], end[:-1]]).astype(np.int64); text = b"".join(enc)
d = torch.from_numpy(emb).cuda()
out = pkg.abi.host_alloc((n * (dim * 17 + 48) + len(text) * 6 + 64,), np.uint8)
with pkg.abi.JsonGpu(0) as jg:
    for _ in range(2):
        doc = jg.sentence_items(text, begin, end, d.data_ptr(), n, dim, out=out)
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        doc = jg.sentence_items(text, begin, end, d.data_ptr(), n, dim, out=out)
    dt = (time.perf_counter() - t0) / reps
t0 = time.perf_counter()
host = pkg.abi.json_sentence_items_ranges(text, begin, end, emb)
dth = time.perf_counter() - t0
print(f"{n} sentences x {dim}: {len(doc) / 1e6:.1f} MB of JSON; GPU formatter {dt * 1e3:.2f} ms = {n / dt / 1e6:.2f} M sentences/s "
      f"({len(doc) / dt / 1e9:.1f} GB/s of text incl. D2H); host formatter ({len(os.sched_getaffinity(0))} cores) {dth * 1e3:.1f} ms = "
      f"{n / dth / 1e6:.2f} M sentences/s; identical: {doc.tobytes() == host.tobytes()}")
