"""GPU: one semantic-search request end to end through the pieces this repo provides (f1 -> path -> f4): the query text is
tokenised natively, embedded (one sentence, CUDA-graph replay), and searched in a resident index of n vectors.
The reference does the same with tokenizers + candle (pad to 512 tokens, chunk of 8) + a Qdrant round trip
(preprocessing_service/src/main.rs:173-298, vector_memory_service/src/main.rs:225-420). Usage: python scripts/query_e2e.py [n]"""
import importlib, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
pkg = importlib.import_module("codename-symbiont_b200")
abi = pkg.abi
from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, processors, trainers
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
shape = pkg.weights.MINILM_L6
corpus = ["the quick brown fox jumps over the lazy dog", "semantic search over sentence embeddings", "how do i reset my password",
          "what is the capital of france", "gpu kernels for transformer encoders"] * 50
tok = Tokenizer(models.WordPiece(unk_token="[UNK]"))
tok.normalizer = normalizers.BertNormalizer(lowercase=True)
tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
tok.train_from_iterator(corpus, trainers.WordPieceTrainer(vocab_size=300, special_tokens=["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"], show_progress=False))
tok.post_processor = processors.TemplateProcessing(single="[CLS] $A [SEP]", special_tokens=[("[CLS]", tok.token_to_id("[CLS]")), ("[SEP]", tok.token_to_id("[SEP]"))])
d = tempfile.mkdtemp(); tok.save(os.path.join(d, "tokenizer.json"))
ing = abi.Ingest(os.path.join(d, "tokenizer.json"))
emb = abi.Embedder.random_init(shape, seed=1)
ix = abi.Index(shape.hidden_size, capacity_hint=n)
g = torch.Generator(device="cuda").manual_seed(0)
for i in range(0, n, 250_000):
    m = min(250_000, n - i)
    v = torch.randn((m, shape.hidden_size), device="cuda", generator=g)
    ix.add_device(v.data_ptr(), m)
query = "how do i reset my password quickly?"
P = 32  # the query is padded to 32 tokens here; the reference pads it to max_position_embeddings
def once():
    t = [time.perf_counter()]
    ids, mask, tt, _ = ing.encode([query], P, threads=1); t.append(time.perf_counter())
    vec = emb.forward(ids, mask, tt); t.append(time.perf_counter())
    hit_ids, scores = ix.search(vec, 10); t.append(time.perf_counter())
    return np.diff(t)
for stage in ("one-stage f32 scan", "two-stage (bf16 copy)"):
  if stage.startswith("two"):
      ix.enable_prescan()
  for _ in range(20): once()
  ts = np.array([once() for _ in range(200)])
  med = np.median(ts, axis=0) * 1e3
  print(f"{stage}: index {n} x {shape.hidden_size}: tokenise {med[0]:.3f} ms, embed {med[1]:.3f} ms, search top-10 {med[2]:.3f} ms, "
        f"total {np.median(ts.sum(1)) * 1e3:.3f} ms (p90 {np.percentile(ts.sum(1), 90) * 1e3:.3f} ms)")
