"""CPU emulation: fp16 vs bf16 MMA operands, with and without the low half of the residual stream.
Decides whether raw max-abs <= 2e-3 (BASELINE gate) is reachable. Not part of the product or the tests.
usage: precision_study_f16.py [minilm-l6|bert-base] """
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()

def rnd(x, fmt):
    if fmt == "f32": return x
    if fmt == "bf16": return x.bfloat16().float()
    if fmt == "f16": return x.half().float()
    if fmt == "bf16x2":  # (hi, lo) pair of bf16
        hi = x.bfloat16().float(); return hi + (x - hi).bfloat16().float()
    if fmt == "f16x2":
        hi = x.half().float(); return hi + (x - hi).half().float()
    if fmt == "f16+e5m2":  # fp16 hi, the remainder as an 8-bit e5m2 float
        hi = x.half().float(); return hi + (x - hi).to(torch.float8_e5m2).float()
    if fmt == "bf16+e5m2":
        hi = x.bfloat16().float(); return hi + (x - hi).to(torch.float8_e5m2).float()
    if fmt == "f16+e4m3s":  # remainder scaled by 2^12 into e4m3
        hi = x.half().float(); return hi + ((x - hi) * 4096).to(torch.float8_e4m3fn).float() / 4096
    raise ValueError(fmt)

def forward(shape, w, ids, mask, tt, op, resid, final=None):
    """op: format of every MMA operand (weights, GEMM inputs, q/k/v, p, ctx, inter); resid: format the LN outputs are
    stored in for the residual add and the pooling."""
    final = final or resid
    H, L, heads = shape.hidden_size, shape.num_hidden_layers, shape.num_attention_heads
    d = H // heads
    W = {k: torch.from_numpy(v) for k, v in w.items()}
    def lin(x, name): return x @ rnd(W[name + ".weight"], op).t() + W[name + ".bias"]
    def ln(x, name): return torch.nn.functional.layer_norm(x, (H,), W[name + ".weight"], W[name + ".bias"], shape.layer_norm_eps)
    ids_t = torch.from_numpy(ids.astype(np.int64)); tt_t = torch.from_numpy(tt.astype(np.int64))
    B, S = ids.shape
    e = W["embeddings.word_embeddings.weight"][ids_t] + W["embeddings.token_type_embeddings.weight"][tt_t] + W["embeddings.position_embeddings.weight"][:S][None]
    x = ln(e, "embeddings.LayerNorm")
    bias = (1.0 - torch.from_numpy(mask.astype(np.float32)))[:, None, None, :] * torch.finfo(torch.float32).min
    for l in range(L):
        p = f"encoder.layer.{l}."
        xr, xg = rnd(x, resid), rnd(x, op)
        q, k, v = (rnd(lin(xg, p + f"attention.self.{n}"), op).view(B, S, heads, d).transpose(1, 2) for n in ("query", "key", "value"))
        sc = q @ k.transpose(-1, -2) / d ** 0.5 + bias
        pe = torch.exp(sc - sc.max(-1, keepdim=True).values)
        ctx = (rnd(pe, op) @ v) / pe.sum(-1, keepdim=True)
        ctx = rnd(ctx.transpose(1, 2).reshape(B, S, H), op)
        a = ln(lin(ctx, p + "attention.output.dense") + xr, p + "attention.output.LayerNorm")
        ar, ag = rnd(a, resid), rnd(a, op)
        i = rnd(torch.nn.functional.gelu(lin(ag, p + "intermediate.dense")), op)
        x = ln(lin(i, p + "output.dense") + ar, p + "output.LayerNorm")
    x = rnd(x, final)
    m = torch.from_numpy(mask.astype(np.float32))[:, :, None]
    return ((x * m).sum(1) / (m.sum(1) + 1e-9)).numpy()

def main():
    skey = sys.argv[1] if len(sys.argv) > 1 else "minilm-l6"
    shape = pkg.weights.SHAPES[skey]
    w = pkg.weights.init_weights(shape, 1234)
    P = 128 if skey == "bert-base" else 512
    lengths = np.array([2, 2, 3, 4, 5, 8, 8, 16, 32, 64, 128] + ([] if P == 128 else [256, 511, 512]))
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, P, seed=42)
    ref = forward(shape, w, ids, mask, tt, "f32", "f32")
    print("ref |x| rms %.3f max %.3f" % (np.sqrt((ref**2).mean()), np.abs(ref).max()))
    for name, op, resid in [("bf16 operands, (hi,lo) bf16 residual  [shipped r1]", "bf16", "bf16x2"),
                            ("bf16 operands, bf16 residual", "bf16", "bf16"),
                            ("fp16 operands, fp32 residual", "f16", "f32"),
                            ("fp16 operands, (hi,lo) fp16 residual", "f16", "f16x2"),
                            ("fp16 operands, fp16 residual (no lo)", "f16", "f16"),
                            ("fp16 weights only", None, None)]:
        if op is None:
            W2 = {k: (rnd(torch.from_numpy(v), "f16").numpy() if k.endswith("weight") and "encoder" in k and "LayerNorm" not in k else v) for k, v in w.items()}
            got = forward(shape, W2, ids, mask, tt, "f32", "f32")
        else:
            got = forward(shape, w, ids, mask, tt, op, resid)
        cos = ((got * ref).sum(1) / np.linalg.norm(got, axis=1) / np.linalg.norm(ref, axis=1)).min()
        err = np.abs(got - ref)
        print(f"{name:52s} cos {cos:.7f} max-abs {err.max():.2e} | by len " + " ".join(f"{n}:{err[i].max():.1e}" for i, n in enumerate(lengths)))
if len(sys.argv) <= 2: main()

def variants():
    """which LayerNorm outputs need the low half when the operands are fp16"""
    skey = sys.argv[1] if len(sys.argv) > 1 else "minilm-l6"
    shape = pkg.weights.SHAPES[skey]
    w = pkg.weights.init_weights(shape, 1234)
    rng = np.random.default_rng(5)
    lengths = np.concatenate([np.full(24, 2), np.full(16, 3), np.full(8, 4), np.array([5, 6, 8, 16, 32, 64, 128])])
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 128, seed=43)
    global forward
    base_forward = forward
    ref = base_forward(shape, w, ids, mask, tt, "f32", "f32")
    def run(op, rx, ra, final):
        # patch: separate formats for x (layer output) and a (post-attention LN)
        H, L, heads = shape.hidden_size, shape.num_hidden_layers, shape.num_attention_heads
        d = H // heads
        W = {k: torch.from_numpy(v) for k, v in w.items()}
        def lin(x, name): return x @ rnd(W[name + ".weight"], op).t() + W[name + ".bias"]
        def ln(x, name): return torch.nn.functional.layer_norm(x, (H,), W[name + ".weight"], W[name + ".bias"], shape.layer_norm_eps)
        ids_t = torch.from_numpy(ids.astype(np.int64)); tt_t = torch.from_numpy(tt.astype(np.int64))
        B, S = ids.shape
        e = W["embeddings.word_embeddings.weight"][ids_t] + W["embeddings.token_type_embeddings.weight"][tt_t] + W["embeddings.position_embeddings.weight"][:S][None]
        x = ln(e, "embeddings.LayerNorm")
        bias = (1.0 - torch.from_numpy(mask.astype(np.float32)))[:, None, None, :] * torch.finfo(torch.float32).min
        for l in range(L):
            p = f"encoder.layer.{l}."
            xr, xg = rnd(x, rx), rnd(x, op)
            q, k, v = (rnd(lin(xg, p + f"attention.self.{n}"), op).view(B, S, heads, d).transpose(1, 2) for n in ("query", "key", "value"))
            sc = q @ k.transpose(-1, -2) / d ** 0.5 + bias
            pe = torch.exp(sc - sc.max(-1, keepdim=True).values)
            ctx = (rnd(pe, op) @ v) / pe.sum(-1, keepdim=True)
            ctx = rnd(ctx.transpose(1, 2).reshape(B, S, H), op)
            a = ln(lin(ctx, p + "attention.output.dense") + xr, p + "attention.output.LayerNorm")
            ar, ag = rnd(a, ra), rnd(a, op)
            i = rnd(torch.nn.functional.gelu(lin(ag, p + "intermediate.dense")), op)
            x = ln(lin(i, p + "output.dense") + ar, p + "output.LayerNorm")
        x = rnd(x, final)
        m = torch.from_numpy(mask.astype(np.float32))[:, :, None]
        return ((x * m).sum(1) / (m.sum(1) + 1e-9)).numpy()
    for name, args in [("f16 op; x f16x2, a f16x2, final f16x2", ("f16", "f16x2", "f16x2", "f16x2")),
                       ("f16 op; x f16,   a f16,   final f16", ("f16", "f16", "f16", "f16")),
                       ("f16 op; x f16,   a f16,   final f32", ("f16", "f16", "f16", "f32")),
                       ("f16 op; x f16x2, a f16,   final f16x2", ("f16", "f16x2", "f16", "f16x2")),
                       ("f16 op; x f16,   a f16x2, final f16x2", ("f16", "f16", "f16x2", "f16x2")),
                       ("f16 op; x, a, final f16+e5m2", ("f16", "f16+e5m2", "f16+e5m2", "f16+e5m2")),
                       ("f16 op; x, a, final f16+e4m3s", ("f16", "f16+e4m3s", "f16+e4m3s", "f16+e4m3s")),
                       ("bf16 op; x, a, final bf16+e5m2", ("bf16", "bf16+e5m2", "bf16+e5m2", "bf16+e5m2")),
                       ("bf16 op; x bf16x2, a bf16x2 [r1]", ("bf16", "bf16x2", "bf16x2", "bf16x2"))]:
        got = run(*args)
        err = np.abs(got - ref).max(1)
        cos = ((got * ref).sum(1) / np.linalg.norm(got, axis=1) / np.linalg.norm(ref, axis=1)).min()
        print(f"{name:44s} cos {cos:.7f} max {err.max():.2e}  len2 {err[:24].max():.2e} len3 {err[24:40].max():.2e} len4 {err[40:48].max():.2e} longer {err[48:].max():.2e}")
if len(sys.argv) > 2 and sys.argv[2] == "variants":
    variants()
