"""GPU diagnostic: is the forward bit-identical run to run, and under sentence permutation? which kernel breaks it?"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gpu_util import dev, dev_i32, bf16_round, host, padded_to_packed, cu_of

shape = pkg.weights.MINILM_L6
lengths = pkg.synth.lengths_uniform(1024, 8, 128, seed=3)
ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 128)
perm = np.random.default_rng(0).permutation(len(lengths))
with pkg.Embedder.random_init(shape, seed=1234) as e:
    a = e.forward(ids, mask, tt); b = e.forward(ids, mask, tt)
    c = e.forward(ids[perm], mask[perm], tt[perm])
    print("env", os.environ.get("B200_EMBED_ATTENTION"), "run-to-run equal", np.array_equal(a, b), "perm equal", np.array_equal(c, a[perm]),
          "max diff", np.abs(c - a[perm]).max(), "rows differing", int((np.abs(c - a[perm]).max(1) > 0).sum()))
    # attention op alone under permutation
    H, heads, d = 384, 12, 32
    rng = np.random.default_rng(5)
    B = len(lengths)
    rows = [bf16_round(rng.standard_normal((n, 3 * H)).astype(np.float32)) for n in lengths]
    def attn(order, dt):
        qkv = np.concatenate([rows[i] for i in order], axis=0)
        cu = cu_of(lengths[order])
        dq, dcu = dev(qkv, torch.bfloat16), dev_i32(cu)
        ctx = torch.zeros((qkv.shape[0], H), dtype=torch.bfloat16, device="cuda")
        e.op_attention(dq.data_ptr(), dcu.data_ptr(), B, qkv.shape[0], 128, dt, ctx.data_ptr())
        out = host(ctx)
        return [out[cu[i]:cu[i + 1]] for i in range(B)]
    for dt in (0, 2):
        x = attn(np.arange(B), dt); y = attn(perm, dt)
        bad = [i for i in range(B) if not np.array_equal(y[i], x[perm[i]])]
        print("attention dtype", dt, "sentences differing under permutation:", len(bad),
              "max", max([np.abs(y[i] - x[perm[i]]).max() for i in bad], default=0.0))
