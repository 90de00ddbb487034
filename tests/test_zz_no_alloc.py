"""GPU: b200_op_linear_noalloc, the entry point for callers that cannot allocate device memory (a process that gave
the whole GPU to its own pool, or a latency-bound loop that cannot take cudaMalloc's device-wide synchronisation).
It must give what b200_op_linear gives -- same kernels, the fp32 temporary in the library's built-in scratch instead
of a cudaMalloc'ed buffer -- on back-to-back calls that reuse that scratch, match the CPU oracle, and refuse sizes
beyond the scratch. Kept in a file of its own that pytest collects last."""
import numpy as np
import pytest
import torch

from gpu_util import dev, round16, T16, host, split_lo8, nan_lo8, LO8_REL
from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["f16", "bf16", "fp32"])
def test_noalloc_linear_matches_allocating_entry_point(pkg, built, mode):
    # hidden 640: no cluster shape owns whole rows of that width, so in the 16-bit modes the LayerNorm is NOT fused
    # into the GEMM and b200_op_linear really parks the fp32 rows in its cudaMalloc'ed temporary (128 ... 1024 in steps
    # the fused epilogue covers - 128, 256, 384, 512, 768, 1024 - would compare against the fused kernel instead);
    # the fp32 mode always uses the temporary
    shape = pkg.weights.BertShape(512, 640, 1, 10, 1280, 64, 2)
    H, I = shape.hidden_size, shape.intermediate_size
    dtype = 1 if mode == "fp32" else 0
    fmt = "f16" if mode == "fp32" else mode
    td = torch.float32 if dtype else T16[fmt]
    rnd = (lambda a: a) if dtype else (lambda a: round16(a, fmt))
    rng = np.random.default_rng(17)
    with pkg.Embedder.random_init(shape, seed=3, operands=fmt) as emb:
        w = emb.export_weights()
        o = Oracle(shape, {k: rnd(v) if v.ndim == 2 else v for k, v in w.items()}, "f32")
        for call, (which, rows) in enumerate([(3, 700), (1, 1500), (3, 64)]):  # consecutive calls share the scratch
            kin = I if which == 3 else H
            x = rnd(rng.standard_normal((rows, kin)).astype(np.float32))
            res, dr_lo = rng.standard_normal((rows, H)).astype(np.float32), None
            if dtype:
                dr = dev(res, td)
            else:
                res, dr, dr_lo = split_lo8(res, fmt)  # the residual as the kernel reads it back: hi + lo
            dx = dev(x, td)
            outs = []
            for fn in (emb.op_linear, emb.op_linear_noalloc):
                y = torch.full((rows, H), float("nan"), dtype=td, device="cuda")
                y_lo = nan_lo8((rows, H))
                fn(0, which, dx.data_ptr(), dr.data_ptr(), rows, dtype, y.data_ptr(),
                   d_residual_lo=0 if dtype else dr_lo.data_ptr(), d_y_lo=0 if dtype else y_lo.data_ptr())
                outs.append((host(y), None if dtype else host(y_lo)))
            (want, want_lo), (got, got_lo) = outs
            assert np.isfinite(got).all()
            # both entry points run the same two kernels on the same inputs: the same bits
            assert np.array_equal(got, want), (call, np.abs(got - want).max())
            if not dtype:
                assert np.array_equal(got_lo, want_lo)
            # and against the CPU oracle, with the tolerances of test_k2_linear_layers
            pre = "attention.output" if which == 1 else "output"
            ref = o.add_layer_norm(o.linear(x, o.layer(0, pre + ".dense.weight"), o.layer(0, pre + ".dense.bias")), res,
                                   o.layer(0, pre + ".LayerNorm.weight"), o.layer(0, pre + ".LayerNorm.bias"),
                                   shape.layer_norm_eps)
            tol = 2e-5 if dtype else 0.006 * np.abs(ref).max() + 1e-3
            assert np.abs(got - ref).max() <= tol, (call, np.abs(got - ref).max())
            if not dtype:
                assert np.abs(got + got_lo - ref).max() <= 2e-3 + LO8_REL[fmt] * np.abs(ref).max()
        # sizes beyond the built-in scratch are refused, and the handle keeps working afterwards
        too_many = pkg.abi.NOALLOC_MAX_ELEMS // H + 1
        big = torch.zeros((too_many, I), dtype=td, device="cuda")
        bigr = torch.zeros((too_many, H), dtype=td, device="cuda")
        with pytest.raises(pkg.B200EmbedError) as e:
            emb.op_linear_noalloc(0, 3, big.data_ptr(), bigr.data_ptr(), too_many, dtype, bigr.data_ptr())
        assert e.value.code == -1 and "scratch" in e.value.message
        with pytest.raises(pkg.B200EmbedError):
            emb.op_linear_noalloc(0, 2, big.data_ptr(), bigr.data_ptr(), 8, dtype, bigr.data_ptr())
