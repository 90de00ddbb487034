"""CPU emulation of where bf16 rounding is applied along the path, to choose which tensors must stay fp32
to meet the BASELINE gate (cos >= 0.9999, max-abs <= 2e-3). Not part of the product or the tests."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()

def r(x, on):
    return x.bfloat16().float() if on else x

def forward(shape, w, ids, mask, tt, R):
    """R: dict of booleans: weights, tables, resid (LN outputs stored bf16 for the residual), gemm_in (A operands bf16),
    qkv, p, ctx, inter, final"""
    H, L, heads = shape.hidden_size, shape.num_hidden_layers, shape.num_attention_heads
    d = H // heads
    W = {k: torch.from_numpy(v) for k, v in w.items()}
    def lin(x, name):
        return x @ r(W[name + ".weight"], R["weights"]).t() + W[name + ".bias"]
    def ln(x, name):
        return torch.nn.functional.layer_norm(x, (H,), W[name + ".weight"], W[name + ".bias"], shape.layer_norm_eps)
    ids_t = torch.from_numpy(ids.astype(np.int64)); tt_t = torch.from_numpy(tt.astype(np.int64))
    B, S = ids.shape
    e = r(W["embeddings.word_embeddings.weight"], R["tables"])[ids_t] + r(W["embeddings.token_type_embeddings.weight"], R["tables"])[tt_t] \
        + r(W["embeddings.position_embeddings.weight"], R["tables"])[:S][None]
    x = ln(e, "embeddings.LayerNorm")
    bias = (1.0 - torch.from_numpy(mask.astype(np.float32)))[:, None, None, :] * torch.finfo(torch.float32).min
    for l in range(L):
        p = f"encoder.layer.{l}."
        xr = r(x, R["resid"])           # what the residual add reads
        xg = r(x, R["gemm_in"])         # what the GEMM reads
        q, k, v = (r(lin(xg, p + f"attention.self.{n}"), R["qkv"]).view(B, S, heads, d).transpose(1, 2) for n in ("query", "key", "value"))
        sc = q @ k.transpose(-1, -2) / d ** 0.5 + bias
        mx = sc.max(-1, keepdim=True).values
        pe = torch.exp(sc - mx)
        ctx = (r(pe, R["p"]) @ v) / pe.sum(-1, keepdim=True)
        ctx = r(ctx.transpose(1, 2).reshape(B, S, H), R["ctx"])
        a = ln(lin(ctx, p + "attention.output.dense") + xr, p + "attention.output.LayerNorm")
        ar, ag = r(a, R["resid"]), r(a, R["gemm_in"])
        i = r(torch.nn.functional.gelu(lin(ag, p + "intermediate.dense")), R["inter"])
        x = ln(lin(i, p + "output.dense") + ar, p + "output.LayerNorm")
    x = r(x, R["final"])
    m = torch.from_numpy(mask.astype(np.float32))[:, :, None]
    return ((x * m).sum(1) / (m.sum(1) + 1e-9)).numpy()

def main():
    skey = sys.argv[1] if len(sys.argv) > 1 else "minilm-l6"
    shape = pkg.weights.SHAPES[skey]
    w = pkg.weights.init_weights(shape, 1234)
    S = min(128, shape.max_position_embeddings)
    lengths = pkg.synth.lengths_uniform(16, 8, S, seed=7); lengths[:4] = S
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, S, seed=42)
    none = dict(weights=False, tables=False, resid=False, gemm_in=False, qkv=False, p=False, ctx=False, inter=False, final=False)
    ref = forward(shape, w, ids, mask, tt, none)
    print("ref |x| rms %.3f max %.3f" % (np.sqrt((ref**2).mean()), np.abs(ref).max()))
    allb = {k: True for k in none}
    cases = {"all bf16 (current)": allb,
             "weights only": dict(none, weights=True),
             "weights+gemm_in+qkv+p+ctx+inter (fp32 resid/tables/final)": dict(allb, resid=False, tables=False, final=False),
             "same but bf16 tables": dict(allb, resid=False, final=False),
             "same but bf16 final": dict(allb, resid=False, tables=False),
             "resid bf16 only": dict(none, resid=True),
             "gemm_in only": dict(none, gemm_in=True),
             "qkv only": dict(none, qkv=True), "p only": dict(none, p=True), "ctx only": dict(none, ctx=True),
             "inter only": dict(none, inter=True), "tables only": dict(none, tables=True), "final only": dict(none, final=True)}
    for name, R in cases.items():
        got = forward(shape, w, ids, mask, tt, R)
        cos = ((got * ref).sum(1) / np.linalg.norm(got, axis=1) / np.linalg.norm(ref, axis=1)).min()
        err = np.abs(got - ref)
        print(f"{name:60s} cos {cos:.7f} max-abs {err.max():.2e} rms {np.sqrt((err**2).mean()):.2e}")

main()


def by_length():
    """max-abs error per sentence length of the shipped design vs the same-weights fp32 result."""
    shape = pkg.weights.MINILM_L6
    w = pkg.weights.init_weights(shape, 1234)
    lengths = np.array([2, 3, 4, 8, 8, 16, 32, 64, 128, 256, 511, 512])
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, 512, seed=42)
    none = dict(weights=False, tables=False, resid=False, gemm_in=False, qkv=False, p=False, ctx=False, inter=False, final=False)
    wonly = forward(shape, w, ids, mask, tt, dict(none, weights=True))
    ref = forward(shape, w, ids, mask, tt, none)
    ship = forward(shape, w, ids, mask, tt, dict(none, weights=True, gemm_in=True, qkv=True, p=True, ctx=True, inter=True))
    for i, n in enumerate(lengths):
        print(f"len {n:4d}: ship vs same-weights {np.abs(ship[i]-wonly[i]).max():.2e}   ship vs fp32 {np.abs(ship[i]-ref[i]).max():.2e}"
              f"   weights-only vs fp32 {np.abs(wonly[i]-ref[i]).max():.2e}   |v|max {np.abs(ref[i]).max():.2f}")

if len(sys.argv) > 2 and sys.argv[2] == "bylen":
    by_length()
