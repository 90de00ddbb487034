"""Makes tests/golden/xlmr_like_tokenizer.json: a Unigram tokenizer.json of the shape XLM-R's has (the tokenizer of the service's
deployed model, main.rs:305), built offline: a SentencePiece unigram model trained on a synthetic multi-script corpus with the
nmt_nfkc normalisation rule (the real precompiled charsmap, 240 KB), wrapped by the `tokenizers` API exactly like the
published file: normalizer Sequence[Precompiled, Replace(" {2,}" -> " ")], pre-tokenizer Sequence[WhitespaceSplit,
Metaspace], Unigram(unk_id 3), TemplateProcessing "<s> $A </s>", specials <s> <pad> </s> <unk> ... <mask>."""
import os, sys, tempfile
import numpy as np
import sentencepiece as spm
from sentencepiece import sentencepiece_model_pb2 as pb
from tokenizers import Tokenizer, Regex, models, normalizers, pre_tokenizers, processors, decoders

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rng = np.random.default_rng(0)
scripts = [(0x61, 0x7a)] * 4 + [(0x41, 0x5a), (0xe0, 0xff), (0x430, 0x44f), (0x3b1, 0x3c9), (0x4e00, 0x4f00), (0x3041, 0x3090),
                                (0xac00, 0xad00), (0x5d0, 0x5ea), (0x627, 0x64a), (0x905, 0x939), (0xe01, 0xe30), (0x30, 0x39)]
punct = list(".,;:!?-()\"'")
lines = []
for i in range(30000):
    lo, hi = scripts[rng.integers(len(scripts))]
    words = []
    for _ in range(rng.integers(3, 14)):
        w = "".join(chr(rng.integers(lo, hi + 1)) for _ in range(rng.integers(1, 9)))
        if rng.random() < 0.15:
            w += punct[rng.integers(len(punct))]
        words.append(w)
    lines.append(" ".join(words))
with tempfile.TemporaryDirectory() as d:
    corpus = os.path.join(d, "corpus.txt")
    open(corpus, "w").write("\n".join(lines))
    prefix = os.path.join(d, "m")
    spm.SentencePieceTrainer.train(input=corpus, model_prefix=prefix, vocab_size=6000, model_type="unigram",
                                   normalization_rule_name="nmt_nfkc", character_coverage=0.9995, pad_id=1, unk_id=3, bos_id=0,
                                   eos_id=2, pad_piece="<pad>", unk_piece="<unk>", bos_piece="<s>", eos_piece="</s>",
                                   minloglevel=2)
    m = pb.ModelProto()
    m.ParseFromString(open(prefix + ".model", "rb").read())
vocab = [("<s>", 0.0), ("<pad>", 0.0), ("</s>", 0.0), ("<unk>", 0.0)] + [(p.piece, p.score) for p in m.pieces[4:]] + [("<mask>", 0.0)]
tok = Tokenizer(models.Unigram(vocab, unk_id=3, byte_fallback=False))
tok.normalizer = normalizers.Sequence([normalizers.Precompiled(m.normalizer_spec.precompiled_charsmap), normalizers.Replace(Regex(" {2,}"), " ")])
tok.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.WhitespaceSplit(), pre_tokenizers.Metaspace(replacement="▁", prepend_scheme="always")])
tok.post_processor = processors.TemplateProcessing(single="<s> $A </s>", pair="<s> $A </s> </s> $B </s>", special_tokens=[("<s>", 0), ("</s>", 2)])
tok.decoder = decoders.Metaspace(replacement="▁", prepend_scheme="always")
tok.add_special_tokens(["<s>", "<pad>", "</s>", "<unk>", "<mask>"])
out = os.path.join(ROOT, "tests", "golden", "xlmr_like_tokenizer.json")
tok.save(out, pretty=False)
print(out, os.path.getsize(out), "bytes,", len(vocab), "pieces")
