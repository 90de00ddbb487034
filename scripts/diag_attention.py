"""GPU diagnostic: tcgen05 attention at scale -- run-to-run determinism, tc vs mma.sync, both vs the oracle."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gpu_util import dev, dev_i32, bf16_round, host, padded_to_packed, cu_of
from oracle.oracle import Oracle

def run(heads, d, lengths, S, seed=0):
    shape = pkg.weights.BertShape(512, heads * d, 1, heads, 512, 512, 2)
    H = heads * d
    rng = np.random.default_rng(seed)
    B = len(lengths)
    q, k, v = (bf16_round(rng.standard_normal((B, S, H)).astype(np.float32)) for _ in range(3))
    mask = (np.arange(S)[None, :] < lengths[:, None]).astype(np.uint32)
    with pkg.Embedder.random_init(shape, seed=1) as e:
        o = Oracle(shape, e.export_weights(), "f32")
        want = padded_to_packed(o.attention(q, k, v, mask, heads), lengths)
        qkv = np.concatenate([padded_to_packed(t, lengths) for t in (q, k, v)], axis=1)
        dq = dev(qkv, torch.bfloat16)
        d_cu = dev_i32(cu_of(lengths))
        outs = {}
        for name, dt in (("tc", 0), ("tc2", 0), ("mma", 2)):
            ctx = torch.full(want.shape, float("nan"), dtype=torch.bfloat16, device="cuda")
            e.op_attention(dq.data_ptr(), d_cu.data_ptr(), B, qkv.shape[0], int(lengths.max()), dt, ctx.data_ptr())
            outs[name] = host(ctx)
    print(f"heads {heads} d {d} B {B} S {S}: tc==tc2 {np.array_equal(outs['tc'], outs['tc2'])}  "
          f"|tc-oracle| {np.abs(outs['tc']-want).max():.4f}  |mma-oracle| {np.abs(outs['mma']-want).max():.4f}  "
          f"|tc-mma| {np.abs(outs['tc']-outs['mma']).max():.4f}  nan {np.isnan(outs['tc']).sum()}")
    bad = np.abs(outs['tc'] - want).max(axis=1)
    cu = cu_of(lengths)
    sent = np.searchsorted(cu, np.nonzero(bad > 0.03)[0], side="right") - 1
    if len(sent):
        print("  bad sentences (first 20):", np.unique(sent)[:20], "lengths", lengths[np.unique(sent)[:20]])

rng = np.random.default_rng(1)
run(12, 32, np.full(64, 128), 128)
run(12, 32, np.full(1024, 128), 128)
run(12, 32, rng.integers(8, 129, size=1024), 128)
run(12, 64, rng.integers(8, 257, size=256), 256)
run(12, 32, rng.integers(1, 513, size=128), 512)
