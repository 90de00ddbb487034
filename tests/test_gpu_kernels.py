"""GPU: every kernel of the path, called through the C ABI (b200_op_*), against the CPU oracle on the
same seeded inputs. The 16-bit kernels are compared with the oracle running on weights and inputs rounded to the
operand format (so only accumulation order and the output rounding differ); fp32 kernels with the oracle as is.
Every test runs once per operand format (bf16 and IEEE fp16: the same kernels compiled twice, csrc/fmt.h)."""
import numpy as np
import pytest
import torch

from conftest import cos_rows
from gpu_util import dev, dev_u32, dev_i32, round16, T16, host, padded_to_packed, cu_of, split_lo8, nan_lo8, LO8_REL
from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu

LENGTHS = [2, 3, 16, 33, 64, 64, 17, 9, 1, 40]

FMT = "bf16"          # operand format of the running parametrisation (set by the fixture below)
td16 = T16["bf16"]  # its torch dtype


def rnd(a):
    return round16(a, FMT)


@pytest.fixture(scope="module", params=["bf16", "f16"], autouse=True)
def fmt(request):
    global FMT, td16
    FMT, td16 = request.param, T16[request.param]
    return request.param


@pytest.fixture(scope="module")
def env(pkg, built, fmt):
    shape = pkg.weights.TINY
    emb = pkg.Embedder.random_init(shape, seed=1234, device=0, operands=fmt)
    assert emb.operands == fmt
    w = emb.export_weights()
    w16 = {k: rnd(v) if v.ndim == 2 else v for k, v in w.items()}
    yield pkg, shape, emb, w, Oracle(shape, w, "f32"), Oracle(shape, w16, "f32")
    emb.close()


def test_native_random_init_matches_numpy_spec(env):
    pkg, shape, emb, w, _, _ = env
    want = pkg.weights.init_weights(shape, 1234)
    assert set(w) == set(want)
    for k in want:
        assert w[k].shape == want[k].shape
        assert np.abs(w[k] - want[k]).max() <= 1e-8, k


def test_safetensors_load_equals_random_init(env, tmp_path):
    pkg, shape, emb, w, _, _ = env
    p = str(tmp_path / "model.safetensors")
    pkg.weights.save_safetensors(p, {"bert." + k: v for k, v in w.items()})  # candle's prefix fallback
    with pkg.Embedder(shape, p, device=0, operands=FMT) as e2:
        w2 = e2.export_weights()
    assert all(np.array_equal(w[k], w2[k]) for k in w)


@pytest.mark.parametrize("dtype", [0, 1])
def test_k1_embed_layernorm(env, dtype):
    pkg, shape, emb, w, o32, o16 = env
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [max(n, 2) for n in LENGTHS], 64, seed=5)
    tt = ((np.arange(64)[None, :] % 3 == 0) & (mask > 0)).astype(np.uint32)
    lengths = mask.sum(1)
    want = padded_to_packed(o32.embeddings(ids, tt), lengths)
    pi, pt, cu = pkg.synth.pack(ids, mask, tt)
    T = int(cu[-1])
    td = torch.float32 if dtype else td16
    out = torch.empty((T, shape.hidden_size), dtype=td, device="cuda")
    out_lo = nan_lo8(out.shape)
    d_pi, d_pt, d_cu = dev_u32(pi), dev_u32(pt), dev_i32(cu)
    emb.op_embed_layernorm(d_pi.data_ptr(), d_pt.data_ptr(), d_cu.data_ptr(), len(lengths), T, dtype, out.data_ptr(),
                           0 if dtype else out_lo.data_ptr())
    got = host(out)
    if dtype:
        assert np.abs(got - want).max() <= 1e-5
    else:  # tables are fp32 in both modes: hi is the bf16 rounding of the exact row, hi + lo restores it
        assert np.abs(got - want).max() <= 2.0 ** -8 * np.abs(want).max()
        assert np.abs(got + host(out_lo) - want).max() <= LO8_REL[FMT] * np.abs(want).max()
    assert cos_rows(got, want).min() >= (0.999999 if dtype else 0.9999)


@pytest.mark.parametrize("m,n,k,block_n,act", [
    (128, 128, 64, 128, -1), (128, 256, 128, 256, -1), (256, 384, 384, 192, -1), (1000, 1152, 384, 192, -1),
    (777, 1536, 384, 256, 0), (300, 512, 128, 128, 0), (513, 256, 1536, 256, -1), (129, 128, 72, 128, 2),
    (4096, 2304, 768, 256, -1), (640, 3072, 768, 256, 1), (1, 128, 64, 128, -1), (20000, 384, 384, 128, -1),
])
def test_gemm_tcgen05_vs_fp32(env, m, n, k, block_n, act):
    """K2 mainloop + bias/activation epilogue vs an fp32 reference of the same op on the same bf16 inputs."""
    pkg = env[0]
    g = torch.Generator(device="cuda").manual_seed(m * 7 + n + k)
    a = (torch.randn((m, k), device="cuda", generator=g) * 1.0).to(td16)
    b = (torch.randn((n, k), device="cuda", generator=g) * 0.05).to(td16)
    bias = torch.randn((n,), device="cuda", generator=g) * 0.1
    c = torch.full((m, n), float("nan"), dtype=td16, device="cuda")
    (pkg.abi.op_gemm_f16 if FMT == "f16" else pkg.abi.op_gemm_bf16)(a.data_ptr(), b.data_ptr(), bias.data_ptr(), m, n, k, act, block_n, c.data_ptr())
    torch.cuda.synchronize()
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = a.float() @ b.float().t() + bias
    if act == 0:
        ref = torch.nn.functional.gelu(ref)
    elif act == 1:
        ref = torch.nn.functional.gelu(ref, approximate="tanh")
    elif act == 2:
        ref = torch.relu(ref)
    err = (c.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert np.isfinite(err) and err <= 0.006 * scale + 1e-3, (err, scale)


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("which", [0, 1, 2, 3])
def test_k2_linear_layers(env, dtype, which):
    """Each Linear of a layer with its fused epilogue: QKV / +res+LN / +GELU / +res+LN (bert.rs)."""
    pkg, shape, emb, w, o32, o16 = env
    o = o32 if dtype else o16
    H, I = shape.hidden_size, shape.intermediate_size
    rows = 333
    rng = np.random.default_rng(which)
    kin = I if which == 3 else H
    x = rng.standard_normal((rows, kin)).astype(np.float32)
    res = rng.standard_normal((rows, H)).astype(np.float32)
    res_hi = res_lo = None
    if not dtype:
        x = rnd(x)
        res, res_hi, res_lo = split_lo8(res, FMT)  # what the kernel reads back as hi + lo
    L = 1
    if which == 0:
        want = np.concatenate([o.linear(x, o.layer(L, f"attention.self.{n}.weight"), o.layer(L, f"attention.self.{n}.bias"))
                               for n in ("query", "key", "value")], axis=1)
    elif which == 1:
        y = o.linear(x, o.layer(L, "attention.output.dense.weight"), o.layer(L, "attention.output.dense.bias"))
        want = o.add_layer_norm(y, res, o.layer(L, "attention.output.LayerNorm.weight"),
                                o.layer(L, "attention.output.LayerNorm.bias"), shape.layer_norm_eps)
    elif which == 2:
        want = o.activation(o.linear(x, o.layer(L, "intermediate.dense.weight"), o.layer(L, "intermediate.dense.bias")), 0)
    else:
        y = o.linear(x, o.layer(L, "output.dense.weight"), o.layer(L, "output.dense.bias"))
        want = o.add_layer_norm(y, res, o.layer(L, "output.LayerNorm.weight"), o.layer(L, "output.LayerNorm.bias"),
                                shape.layer_norm_eps)
    td = torch.float32 if dtype else td16
    dx, dr, dr_lo = dev(x, td), (dev(res, td) if dtype else res_hi), res_lo  # residual = hi + lo in 16-bit mode
    out = torch.full(want.shape, float("nan"), dtype=td, device="cuda")
    out_lo = nan_lo8(want.shape)
    ln = which in (1, 3)
    emb.op_linear(L, which, dx.data_ptr(), dr.data_ptr(), rows, dtype, out.data_ptr(),
                  d_residual_lo=0 if dtype else dr_lo.data_ptr(), d_y_lo=out_lo.data_ptr() if (ln and not dtype) else 0)
    got = host(out)
    tol = 2e-5 if dtype else 0.006 * np.abs(want).max() + 1e-3
    assert np.abs(got - want).max() <= tol, np.abs(got - want).max()
    if ln and not dtype:  # hi + lo carries 3 more bits of the LayerNorm output than hi; the 16-bit GEMM inputs limit it
        assert np.abs(got + host(out_lo) - want).max() <= 2e-3 + LO8_REL[FMT] * np.abs(want).max()


@pytest.mark.parametrize("H,heads,I,rows", [(128, 4, 512, 333), (256, 8, 1024, 1000), (384, 12, 1536, 3000)])
def test_k2f_fused_ffn(pkg, built, H, heads, I, rows):
    """intermediate.dense + GELU + output.dense + residual + LayerNorm in one kernel (bert.rs BertIntermediate +
    BertOutput): the intermediate never leaves the SM. Oracle on the same bf16-rounded weights and input; the
    intermediate is rounded to bf16 on the GPU (it is the second GEMM's operand), hence the tolerance."""
    shape = pkg.weights.BertShape(512, H, 1, heads, I, 64, 2)
    rng = np.random.default_rng(H)
    x, dx, dx_lo = split_lo8(rng.standard_normal((rows, H)).astype(np.float32), FMT)
    x_hi = host(dx)
    with pkg.Embedder.random_init(shape, seed=4, operands=FMT) as emb:
        w = emb.export_weights()
        o = Oracle(shape, {k: rnd(v) if v.ndim == 2 else v for k, v in w.items()}, "f32")
        inter = o.activation(o.linear(x_hi, o.layer(0, "intermediate.dense.weight"), o.layer(0, "intermediate.dense.bias")), 0)
        y = o.linear(rnd(inter), o.layer(0, "output.dense.weight"), o.layer(0, "output.dense.bias"))
        want = o.add_layer_norm(y, x, o.layer(0, "output.LayerNorm.weight"), o.layer(0, "output.LayerNorm.bias"),
                                shape.layer_norm_eps)
        out = torch.full((rows, H), float("nan"), dtype=td16, device="cuda")
        out_lo = nan_lo8((rows, H))
        emb.op_ffn(0, dx.data_ptr(), dx_lo.data_ptr(), rows, out.data_ptr(), out_lo.data_ptr())
        got, got_lo = host(out), host(out_lo)
    assert np.isfinite(got).all() and np.isfinite(got_lo).all()
    assert np.abs(got - want).max() <= 0.006 * np.abs(want).max() + 1e-3
    assert np.abs(got + got_lo - want).max() <= 2e-3 + LO8_REL[FMT] * np.abs(want).max()


@pytest.mark.parametrize("which", [1, 3])
@pytest.mark.parametrize("H,heads,I", [(256, 8, 512), (384, 12, 1536), (512, 8, 1024), (768, 12, 3072), (640, 10, 1280)])
def test_k2_layernorm_epilogue_across_hidden_sizes(pkg, built, H, heads, I, which):
    """The fused residual + LayerNorm epilogue for every row width: one CTA (128, 256), a cluster of 2 CTAs
    (384, 512) or of 4 (768, 1024) exchanging row statistics through distributed shared memory; 640 is not
    fusable and takes the GEMM + K4 route. Rows deliberately not a multiple of 128."""
    shape = pkg.weights.BertShape(512, H, 1, heads, I, 64, 2)
    rows = 1000
    rng = np.random.default_rng(H + which)
    kin = I if which == 3 else H
    x = rnd(rng.standard_normal((rows, kin)).astype(np.float32))
    res, dr, dr_lo = split_lo8(rng.standard_normal((rows, H)).astype(np.float32) * 2, FMT)
    with pkg.Embedder.random_init(shape, seed=2, operands=FMT) as emb:
        w = emb.export_weights()
        o = Oracle(shape, {k: rnd(v) if v.ndim == 2 else v for k, v in w.items()}, "f32")
        pre = "attention.output" if which == 1 else "output"
        y = o.linear(x, o.layer(0, pre + ".dense.weight"), o.layer(0, pre + ".dense.bias"))
        want = o.add_layer_norm(y, res, o.layer(0, pre + ".LayerNorm.weight"), o.layer(0, pre + ".LayerNorm.bias"),
                                shape.layer_norm_eps)
        dx = dev(x, td16)
        out = torch.full(want.shape, float("nan"), dtype=td16, device="cuda")
        out_lo = nan_lo8(want.shape)
        emb.op_linear(0, which, dx.data_ptr(), dr.data_ptr(), rows, 0, out.data_ptr(), d_residual_lo=dr_lo.data_ptr(),
                      d_y_lo=out_lo.data_ptr())
        got, got_lo = host(out), host(out_lo)
    assert np.isfinite(got).all() and np.isfinite(got_lo).all()
    assert np.abs(got - want).max() <= 0.006 * np.abs(want).max() + 1e-3
    assert np.abs(got + got_lo - want).max() <= 2e-3 + LO8_REL[FMT] * np.abs(want).max()


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("heads,d", [(4, 32), (2, 64)])
def test_k3_attention(env, pkg, dtype, heads, d):
    """Varlen attention vs the oracle's padded additive-mask softmax; ragged lengths incl. 1 and > 64."""
    shape = pkg.weights.BertShape(512, heads * d, 1, heads, 512, 256, 2)
    lengths = np.array([1, 2, 63, 64, 65, 128, 200, 17, 129, 256, 255])
    S, H = 256, heads * d
    rng = np.random.default_rng(11)
    q, k, v = (rng.standard_normal((len(lengths), S, H)).astype(np.float32) for _ in range(3))
    if dtype != 1:
        q, k, v = rnd(q), rnd(k), rnd(v)
    mask = (np.arange(S)[None, :] < lengths[:, None]).astype(np.uint32)
    with pkg.Embedder.random_init(shape, seed=1, operands=FMT) as e:
        o = Oracle(shape, e.export_weights(), "f32")
        want = padded_to_packed(o.attention(q, k, v, mask, heads), lengths)
        qkv = np.concatenate([padded_to_packed(t, lengths) for t in (q, k, v)], axis=1)
        td = torch.float32 if dtype == 1 else td16
        dq = dev(qkv, td)
        ctx = torch.full(want.shape, float("nan"), dtype=td, device="cuda")
        d_cu = dev_i32(cu_of(lengths))
        e.op_attention(dq.data_ptr(), d_cu.data_ptr(), len(lengths), qkv.shape[0], int(lengths.max()), dtype,
                       ctx.data_ptr())
        got = host(ctx)
    tol = 2e-5 if dtype == 1 else 0.02  # bf16 P and output rounding on |ctx| <~ 3
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= tol, np.abs(got - want).max()
    assert cos_rows(got, want).min() >= (0.999999 if dtype == 1 else 0.9998)


@pytest.mark.parametrize("heads,d", [(4, 32), (2, 64)])
def test_k3_attention_rescales_when_the_row_maximum_jumps(env, pkg, heads, d):
    """K3 keeps an item's output in one tensor-memory accumulator and lets a unit of 64 keys reuse the row maximum of the
    units before it unless the new maximum is more than 2^8 above; then the accumulator is rescaled in place. Keys whose
    scores grow from unit to unit (and keys that jump only for some rows and some heads) take that path in every unit;
    compared with the oracle's softmax like the plain test."""
    shape = pkg.weights.BertShape(512, heads * d, 1, heads, 512, 512, 2)
    lengths = np.array([512, 300, 129, 65, 511, 64, 200])
    S, H = 512, heads * d
    rng = np.random.default_rng(23)
    q, k, v = (rng.standard_normal((len(lengths), S, H)).astype(np.float32) for _ in range(3))
    # the keys of unit u are scaled by 1 + 0.9 u (scores of later units dominate more and more), one key per unit is an
    # outlier aligned with a few query rows only, and one head is left alone
    unit = np.arange(S) // 64
    k *= (1.0 + 0.9 * unit)[None, :, None]
    for b in range(len(lengths)):
        for u in range(1, 8):
            j = u * 64 + 7 * b % 64
            rows = rng.integers(0, S, size=9)
            k[b, j, :H - d] = 3.0 * q[b, rows[0], :H - d]
    q, k, v = rnd(q), rnd(k), rnd(v)
    mask = (np.arange(S)[None, :] < lengths[:, None]).astype(np.uint32)
    with pkg.Embedder.random_init(shape, seed=1, operands=FMT) as e:
        o = Oracle(shape, e.export_weights(), "f32")
        want = padded_to_packed(o.attention(q, k, v, mask, heads), lengths)
        qkv = np.concatenate([padded_to_packed(t, lengths) for t in (q, k, v)], axis=1)
        dq = dev(qkv, td16)
        ctx = torch.full(want.shape, float("nan"), dtype=td16, device="cuda")
        d_cu = dev_i32(cu_of(lengths))
        e.op_attention(dq.data_ptr(), d_cu.data_ptr(), len(lengths), qkv.shape[0], int(lengths.max()), 0, ctx.data_ptr())
        got = host(ctx)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 0.03, np.abs(got - want).max()
    assert cos_rows(got, want).min() >= 0.9998


@pytest.mark.parametrize("lens", ["mixed", "tiny"])
@pytest.mark.parametrize("heads", [4, 8, 12])
def test_k23_qkv_attention_fused(pkg, built, heads, lens):
    """Q/K/V projection + attention of whole sentences in one kernel (bert.rs BertSelfAttention::forward) vs the
    oracle's Linear + padded additive-mask softmax on the same bf16-rounded weights and input; q, k, v are rounded to
    bf16 on the GPU (they are the attention MMAs' operands). Ragged lengths 1 .. 128 incl. both sides of 64."""
    d = 32
    H = heads * d
    shape = pkg.weights.BertShape(512, H, 1, heads, 4 * H, 128, 2)
    # short sentences share a 128-row tile (block-diagonal attention); "tiny": up to 128 sentences in one tile, tiles
    # cut at the 32-sentence segments of the packing kernel
    lengths = (np.array([1, 2, 63, 64, 65, 128, 100, 17, 127, 128, 33, 5] * 3) if lens == "mixed"
               else np.random.default_rng(7).integers(1, 4, size=301))
    S = 128
    rng = np.random.default_rng(heads)
    x = rnd(rng.standard_normal((len(lengths), S, H)).astype(np.float32))
    mask = (np.arange(S)[None, :] < lengths[:, None]).astype(np.uint32)
    with pkg.Embedder.random_init(shape, seed=5, operands=FMT) as e:
        w = e.export_weights()
        o = Oracle(shape, {k: rnd(v) if v.ndim == 2 else v for k, v in w.items()}, "f32")
        pre = "attention.self."
        q, k, v = (rnd(o.linear(x.reshape(-1, H), o.layer(0, pre + n + ".weight"), o.layer(0, pre + n + ".bias"))
                              .reshape(len(lengths), S, H)) for n in ("query", "key", "value"))
        want = padded_to_packed(o.attention(q, k, v, mask, heads), lengths)
        xp = padded_to_packed(x, lengths)
        dx = dev(xp, td16)
        ctx = torch.full(want.shape, float("nan"), dtype=td16, device="cuda")
        d_cu = dev_i32(cu_of(lengths))
        e.op_qkv_attention(0, dx.data_ptr(), d_cu.data_ptr(), len(lengths), xp.shape[0], S, ctx.data_ptr())
        got = host(ctx)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 0.02 * max(1.0, np.abs(want).max()), np.abs(got - want).max()
    assert cos_rows(got, want).min() >= 0.9998


@pytest.mark.parametrize("mode", ["bf16", "f32y", "fp32"])
def test_k4_add_layernorm(env, mode):
    pkg, shape, emb, w, o32, _ = env
    rows, H = 1001, shape.hidden_size
    rng = np.random.default_rng(3)
    y = rng.standard_normal((rows, H)).astype(np.float32) * 2
    r = rng.standard_normal((rows, H)).astype(np.float32)
    d_r = d_r_lo = None
    if mode != "fp32":
        r, d_r, d_r_lo = split_lo8(r, FMT)
    if mode == "bf16":
        y = rnd(y)
    want = o32.add_layer_norm(y, r, o32.layer(0, "output.LayerNorm.weight"), o32.layer(0, "output.LayerNorm.bias"),
                              shape.layer_norm_eps)
    td = torch.float32 if mode == "fp32" else td16
    dy = dev(y, td16 if mode == "bf16" else torch.float32)
    out = torch.empty((rows, H), dtype=td, device="cuda")
    out_lo = nan_lo8((rows, H))
    if mode == "fp32":
        d_r = dev(r, td)
    emb.op_add_layernorm(0, 3, dy.data_ptr(), mode != "bf16", d_r.data_ptr(), rows, int(mode == "fp32"),
                         out.data_ptr(), d_residual_lo=0 if mode == "fp32" else d_r_lo.data_ptr(),
                         d_out_lo=0 if mode == "fp32" else out_lo.data_ptr())
    got = host(out)
    assert np.abs(got - want).max() <= (1e-5 if mode == "fp32" else 0.02)
    if mode != "fp32":
        assert np.abs(got + host(out_lo) - want).max() <= 1e-5 + LO8_REL[FMT] * np.abs(want).max()


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("normalize", [0, 1])
def test_k5_mean_pool(env, dtype, normalize):
    """embedding_generator.rs:201-207 (+ the north_star L2 step)."""
    pkg, shape, emb, w, o32, _ = env
    lengths = np.array(LENGTHS + [64] * 7)
    S, H = 64, shape.hidden_size
    rng = np.random.default_rng(9)
    hid = rng.standard_normal((len(lengths), S, H)).astype(np.float32)
    mask = (np.arange(S)[None, :] < lengths[:, None]).astype(np.uint32)
    d_lo = None
    if not dtype:  # the rows as the kernel reads them back: hi + lo
        flat, d_hi_flat, d_lo_flat = split_lo8(hid.reshape(-1, H), FMT)
        hid = flat.reshape(hid.shape)
    want = o32.mean_pool(hid, mask)
    if normalize:
        want = o32.l2_normalize(want)
    td = torch.float32 if dtype else td16
    out = torch.empty((len(lengths), H), dtype=torch.float32, device="cuda")
    packed = padded_to_packed(hid, lengths)
    if dtype:
        d_h = dev(packed, td)
    else:  # the packed rows of the pair split above
        idx = torch.from_numpy(np.concatenate([b * S + np.arange(n) for b, n in enumerate(lengths)])).cuda()
        d_h, d_lo = d_hi_flat[idx].contiguous(), d_lo_flat.view(torch.uint8)[idx].contiguous()
    d_cu = dev_i32(cu_of(lengths))
    emb.op_mean_pool(d_h.data_ptr(), d_cu.data_ptr(), len(lengths), dtype, normalize, out.data_ptr(),
                     d_hidden_lo=0 if dtype else d_lo.data_ptr())
    got = host(out)
    assert np.abs(got - want).max() <= (2e-6 if dtype else 3e-5)  # 16-bit: the oracle pools the same hi + lo rows


def test_k2_layernorm_gemm_refuses_operands_outside_this_gpus_memory(pkg, built):
    """The LayerNorm-fused GEMM multicasts its A tile inside a cluster, and multicast loads read only plain device memory
    of the current GPU: an operand in pinned host memory (a stand-in for any mapping that is not a local allocation) is
    refused with an error before anything is launched; the same call with a device operand works."""
    shape = pkg.weights.BertShape(512, 384, 1, 12, 1536, 64, 2)
    rows = 256
    with pkg.Embedder.random_init(shape, seed=2, operands=FMT) as emb:
        x_host = pkg.abi.host_alloc((rows, 384), np.uint16)
        x_host[:] = 0
        res = torch.zeros((rows, 384), dtype=td16, device="cuda")
        out, out_lo = torch.empty_like(res), torch.empty_like(res)
        with pytest.raises(pkg.abi.B200EmbedError):
            emb.op_linear(0, 1, x_host.ctypes.data, res.data_ptr(), rows, 0, out.data_ptr(), d_residual_lo=res.data_ptr(),
                          d_y_lo=out_lo.data_ptr())
        x_dev = torch.zeros((rows, 384), dtype=td16, device="cuda")
        emb.op_linear(0, 1, x_dev.data_ptr(), res.data_ptr(), rows, 0, out.data_ptr(), d_residual_lo=res.data_ptr(),
                      d_y_lo=out_lo.data_ptr())
        torch.cuda.synchronize()
        assert torch.isfinite(out.float()).all()
        pkg.abi.host_free(x_host)
