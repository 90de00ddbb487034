"""GPU: host enqueue time vs device time of one small call (b200_embed_forward_device on a caller stream)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
pkg = load_package()
shape = pkg.weights.MINILM_L6
with pkg.Embedder.random_init(shape, seed=1234) as emb:
    for batch, seq in ((1, 32), (8, 128)):
        ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, [seq] * batch, seq)
        d = [torch.from_numpy(a.astype(np.int64)).to(torch.int32).cuda() for a in (ids, mask, tt)]
        out = torch.empty((batch, shape.hidden_size), dtype=torch.float32, device="cuda")
        s = torch.cuda.Stream()
        f = lambda: emb.forward_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), batch, seq, out.data_ptr(), stream=s.cuda_stream)
        for _ in range(20): f()
        s.synchronize()
        enq, tot, dev = [], [], []
        for _ in range(100):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record(s); f(); e1.record(s)
            t1 = time.perf_counter()
            s.synchronize()
            t2 = time.perf_counter()
            enq.append(t1 - t0); tot.append(t2 - t0); dev.append(e0.elapsed_time(e1))
        print(f"batch {batch} x seq {seq}: host enqueue {np.median(enq)*1e3:.3f} ms, device {np.median(dev):.3f} ms, total {np.median(tot)*1e3:.3f} ms")
