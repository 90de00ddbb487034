"""GPU: b200_op_linear_noalloc, the entry point for callers that cannot allocate device memory (a process that gave
the whole GPU to its own pool, or a latency-bound loop that cannot take cudaMalloc's device-wide synchronisation).
It must give what b200_op_linear gives -- same kernels, the fp32 temporary in the library's built-in scratch instead
of a cudaMalloc'ed buffer -- on back-to-back calls that reuse that scratch, and refuse sizes beyond the scratch.
Kept in a file of its own that pytest collects last."""
import numpy as np
import pytest
import torch

from gpu_util import dev, bf16_round, host

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [0, 1])
def test_noalloc_linear_matches_allocating_entry_point(pkg, built, dtype):
    # hidden 512: the LayerNorm is not fused into the GEMM, so b200_op_linear really uses its temporary
    shape = pkg.weights.BertShape(512, 512, 1, 8, 1024, 64, 2)
    H, I = shape.hidden_size, shape.intermediate_size
    td = torch.float32 if dtype else torch.bfloat16
    rng = np.random.default_rng(17)
    with pkg.Embedder.random_init(shape, seed=3) as emb:
        for call, (which, rows) in enumerate([(3, 700), (1, 1500), (3, 64)]):  # consecutive calls share the scratch
            kin = I if which == 3 else H
            x = rng.standard_normal((rows, kin)).astype(np.float32)
            res = rng.standard_normal((rows, H)).astype(np.float32)
            if not dtype:
                x = bf16_round(x)
            dx, dr, dr_lo = dev(x, td), dev(res, td), dev(res - bf16_round(res), td)
            outs = []
            for fn in (emb.op_linear, emb.op_linear_noalloc):
                y = torch.full((rows, H), float("nan"), dtype=td, device="cuda")
                y_lo = torch.full((rows, H), float("nan"), dtype=td, device="cuda")
                fn(0, which, dx.data_ptr(), dr.data_ptr(), rows, dtype, y.data_ptr(),
                   d_residual_lo=0 if dtype else dr_lo.data_ptr(), d_y_lo=0 if dtype else y_lo.data_ptr())
                outs.append((host(y), None if dtype else host(y_lo)))
            (want, want_lo), (got, got_lo) = outs
            assert np.isfinite(got).all()
            tol = 2e-5 if dtype else 0.006 * np.abs(want).max() + 1e-3  # the tolerances of test_k2_linear_layers
            assert np.abs(got - want).max() <= tol, (call, np.abs(got - want).max())
            if not dtype:
                assert np.abs((got + got_lo) - (want + want_lo)).max() <= 2e-3
        # sizes beyond the built-in scratch are refused, and the handle keeps working afterwards
        too_many = pkg.abi.NOALLOC_MAX_ELEMS // H + 1
        big = torch.zeros((too_many, I), dtype=td, device="cuda")
        bigr = torch.zeros((too_many, H), dtype=td, device="cuda")
        with pytest.raises(pkg.B200EmbedError) as e:
            emb.op_linear_noalloc(0, 3, big.data_ptr(), bigr.data_ptr(), too_many, dtype, bigr.data_ptr())
        assert e.value.code == -1 and "scratch" in e.value.message
        with pytest.raises(pkg.B200EmbedError):
            emb.op_linear_noalloc(0, 2, big.data_ptr(), bigr.data_ptr(), 8, dtype, bigr.data_ptr())
