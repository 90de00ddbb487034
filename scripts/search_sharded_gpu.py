"""GPUs: the row-sharded search (f4 across ranks) on real devices. Launch with
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 scripts/search_sharded_gpu.py
Every rank holds n/N vectors in its own b200_index, searches them, all_gathers the [q, k] lists over NCCL and merges;
rank 0 checks the result against the numpy oracle over all the vectors and prints the time per search."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from __graft_entry__ import load_package
from oracle import search_oracle
pkg = load_package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl")
n, dim, k, q = 400_000, 384, 10, 3
rng = np.random.default_rng(0)  # every rank draws the same data and keeps its slice
vectors = rng.standard_normal((n, dim)).astype(np.float32)
queries = rng.standard_normal((q, dim)).astype(np.float32)
lo, hi = n * rank // world, n * (rank + 1) // world
ix = pkg.abi.Index(dim, capacity_hint=hi - lo, device=local)
ix.add(vectors[lo:hi])
ids, sc = pkg.sharding.search_sharded(ix.search, lo, queries, k)
dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    ids, sc = pkg.sharding.search_sharded(ix.search, lo, queries, k)
dt = (time.perf_counter() - t0) / 20
if rank == 0:
    want_ids, want_sc = search_oracle.search(vectors, queries, k)
    same = np.array_equal(ids, want_ids)
    err = float(np.abs(sc - want_sc).max())
    print(f"world {world}: {n} x {dim} sharded, top-{k} of {q} queries: ids equal oracle: {same}, max score error {err:.2e}, {dt * 1e3:.2f} ms per search")
    assert err <= 2e-6 and (same or err <= 2e-6)
dist.destroy_process_group()
