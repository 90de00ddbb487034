"""GPU: throughput on a corpus-like length distribution (clip(round(LogNormal(ln 24, 0.6)), 4, 128)),
sentences padded to 128 by the tokenizer, device-resident inputs."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
B, S = int(sys.argv[1]) if len(sys.argv) > 1 else 16384, 128
lengths = pkg.synth.lengths_corpus(B)
ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, S)
with pkg.Embedder.random_init(shape, seed=1234) as emb:
    d = [torch.from_numpy(a.astype(np.int64)).to(torch.int32).cuda() for a in (ids, mask, tt)]
    out = torch.empty((B, shape.hidden_size), dtype=torch.float32, device="cuda")
    s = torch.cuda.Stream()
    f = lambda: emb.forward_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), B, S, out.data_ptr(), stream=s.cuda_stream)
    for _ in range(3): f()
    s.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for _ in range(10): f()
    e1.record(s); s.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"corpus lengths (mean {lengths.mean():.1f}, {int(lengths.sum())} tokens), {B} sentences: {ms:.3f} ms -> {B / ms * 1e3:,.0f} sentences/s "
          f"[QKV_ATTN={os.environ.get('B200_EMBED_QKV_ATTN', 'auto')}]")
