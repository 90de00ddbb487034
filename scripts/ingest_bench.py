"""Throughput of the ingest front-end (f1): sentence split + WordPiece encode of an English-like corpus, against the
`tokenizers` library's encode_batch (the reference's own tokenizer, all cores through rayon) on the same sentences.
Usage: python scripts/ingest_bench.py [n_sentences] [seq_len]"""
import importlib, os, random, sys, tempfile, time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

abi = importlib.import_module("codename-symbiont_b200.abi")
from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, processors, trainers

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
P = int(sys.argv[2]) if len(sys.argv) > 2 else 128
rng = random.Random(3)
syll = ["ta", "ri", "mo", "ne", "ka", "su", "lo", "pe", "di", "vu", "an", "er", "in", "st", "ch", "th", "ou", "qu", "ment", "tion", "ing", "ed"]
lex = ["".join(rng.choice(syll) for _ in range(rng.randint(1, 4))) for _ in range(20000)]
weights = [1.0 / (i + 1) for i in range(len(lex))]  # Zipf-like
def sentence():
    k = max(3, int(rng.gauss(18, 7)))
    w = rng.choices(lex, weights, k=k)
    w[0] = w[0].capitalize()
    if rng.random() < 0.3: w[rng.randrange(k)] += ","
    return " ".join(w) + rng.choice([".", ".", ".", "?", "!"])
sents = [sentence() for _ in range(20000)]
doc = "  ".join(rng.choice(sents) for _ in range(n))

tok = Tokenizer(models.WordPiece(unk_token="[UNK]"))
tok.normalizer = normalizers.BertNormalizer(lowercase=True)
tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
tok.train_from_iterator(sents, trainers.WordPieceTrainer(vocab_size=8000, special_tokens=["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"], show_progress=False))
tok.post_processor = processors.TemplateProcessing(single="[CLS] $A [SEP]", pair="[CLS] $A [SEP] $B:1 [SEP]:1",
                                                   special_tokens=[("[CLS]", tok.token_to_id("[CLS]")), ("[SEP]", tok.token_to_id("[SEP]"))])
path = os.path.join(tempfile.mkdtemp(), "tokenizer.json")
tok.save(path)
tok.enable_padding(direction="right", pad_id=tok.token_to_id("[PAD]"), pad_type_id=0, pad_token="[PAD]", length=P)
tok.enable_truncation(max_length=P, stride=0, strategy="longest_first", direction="right")
ing = abi.Ingest(path)

t0 = time.perf_counter()
cleaned, begin, end = abi.split_sentences(doc)
t1 = time.perf_counter()
out = tuple(np.empty((len(begin), P), dtype=np.uint32) for _ in range(3))
ing.encode_ranges(cleaned, begin, end, P, 0, out)  # warm
best = {}
for threads in (1, 0):
    ts = []
    for _ in range(3):
        t2 = time.perf_counter()
        ids, mask, tt, lengths = ing.encode_ranges(cleaned, begin, end, P, threads, out)
        ts.append(time.perf_counter() - t2)
    best[threads] = min(ts)
strs = [cleaned[b:e].decode() for b, e in zip(begin, end)]
t3 = time.perf_counter()
enc = tok.encode_batch(strs, add_special_tokens=True)
t4 = time.perf_counter()
ref_ids = np.array([e.ids for e in enc], dtype=np.uint32)
print(f"sentences {len(begin)}  bytes {len(doc)}  tokens/sentence {lengths.mean():.1f}  cores {os.cpu_count()}")
print(f"split      {len(doc) / (t1 - t0) / 1e6:8.1f} MB/s   ({(t1 - t0) * 1e3:.1f} ms)")
print(f"encode 1 thread   {len(begin) / best[1]:10.0f} sentences/s")
print(f"encode all cores  {len(begin) / best[0]:10.0f} sentences/s")
print(f"tokenizers encode_batch (rayon, all cores) {len(begin) / (t4 - t3):10.0f} sentences/s   ids identical: {np.array_equal(ref_ids, ids)}")
