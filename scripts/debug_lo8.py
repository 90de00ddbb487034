"""GPU debug: where do hi + lo outputs of the LayerNorm epilogue differ from the oracle (by column / row position)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_package
pkg = load_package()
from gpu_util import dev, round16, T16, host, split_lo8, nan_lo8
from oracle.oracle import Oracle
fmt = sys.argv[1] if len(sys.argv) > 1 else "bf16"
H = int(sys.argv[2]) if len(sys.argv) > 2 else 128
shape = pkg.weights.BertShape(512, H, 1, H // 32, 4 * H, 64, 2)
rows = 333
rng = np.random.default_rng(1)
rnd = lambda a: round16(a, fmt)
x = rnd(rng.standard_normal((rows, H)).astype(np.float32))
res, res_lo = split_lo8(rng.standard_normal((rows, H)).astype(np.float32), fmt)
with pkg.Embedder.random_init(shape, seed=2, operands=fmt) as emb:
    w = emb.export_weights()
    o = Oracle(shape, {k: rnd(v) if v.ndim == 2 else v for k, v in w.items()}, "f32")
    y = o.linear(x, o.layer(0, "attention.output.dense.weight"), o.layer(0, "attention.output.dense.bias"))
    for name, r_in, lo_in in (("with lo", res, res_lo), ("zero lo", rnd(res), torch.zeros_like(res_lo))):
        want = o.add_layer_norm(y, r_in, o.layer(0, "attention.output.LayerNorm.weight"), o.layer(0, "attention.output.LayerNorm.bias"), shape.layer_norm_eps)
        td = T16[fmt]
        out = torch.full(want.shape, float("nan"), dtype=td, device="cuda"); out_lo = nan_lo8(want.shape)
        dx, dr = dev(x, td), dev(r_in, td)
        emb.op_linear(0, 1, dx.data_ptr(), dr.data_ptr(), rows, 0, out.data_ptr(), d_residual_lo=lo_in.data_ptr(), d_y_lo=out_lo.data_ptr())
        got, got_lo = host(out), host(out_lo)
        e_hi = np.abs(got - want); e = np.abs(got + got_lo - want)
        # the low half the host would compute from the kernel's own hi
        ideal_lo = torch.from_numpy(want - got).to(torch.float8_e5m2).float().numpy()
        mism = (ideal_lo != got_lo)
        print(f"{fmt} H={H} {name}: hi err {e_hi.max():.3e}, hi+lo err {e.max():.3e}, lo mismatches vs e5m2(want - hi): {mism.mean():.3f}")
        bad = e > 2e-3
        print("   bad elements", bad.sum(), "by col%32", np.bincount(np.nonzero(bad)[1] % 32, minlength=32).tolist())
        print("   by row%32", np.bincount(np.nonzero(bad)[0] % 32, minlength=32).tolist())
