"""BASELINE config 5: a synthetic corpus of N sentences of TEXT ingested over all GPUs of the box in one process
(b200_corpus_*: tokeniser pool -> length-sorted micro-batches -> one queue -> a worker per GPU, two requests in flight each).

  python scripts/ingest_corpus.py --gpus 8 --sentences 10000000 [--json] [--check 32768] [--start-shard K] [--progress FILE]

The corpus: sentences whose token counts follow SURVEY section 8(d)'s c5 law, clip(round(LogNormal(ln 24, 0.6)), 4, 128), written
as text over a synthetic WordPiece vocabulary of 30 522 entries (every word one token, "." ends the sentence, [CLS] / [SEP]
added by the tokenizer), joined into shards of --shard sentences of raw text. Generating text in Python is slow, so
--base-shards distinct shards are generated BEFORE the clock starts and the corpus cycles through them; everything from the raw
text on is inside the timed region, per shard: the reference's splitter (main.rs:28-63, b200_ingest_split_sentences), the
tokeniser, the GPUs, the scatter into input order, and with --json the TextWithEmbeddingsMessage payload (main.rs:138).
Restart (checkpoint row): shards are deterministic functions of their index, so a run resumes with
--start-shard K (the number of the first shard still to do; --progress FILE keeps the last finished one up to date).
Prints one JSON line."""
import argparse, json, os, sys, threading, time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()


def make_tokenizer(path, vocab_words=30522, seed=5):
    """A BERT WordPiece tokenizer.json over a synthetic vocabulary (no network): specials, punctuation, random lowercase words."""
    from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, processors
    rng = np.random.default_rng(seed)
    words, seen = [], set()
    letters = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz", dtype=np.uint8)
    while len(words) < vocab_words - 8:
        n = int(rng.integers(2, 10))
        w = bytes(rng.choice(letters, size=n)).decode()
        if w not in seen:
            seen.add(w)
            words.append(w)
    vocab = {w: i for i, w in enumerate(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", ".", "!", "?"] + words)}
    tok = Tokenizer(models.WordPiece(vocab, unk_token="[UNK]"))
    tok.normalizer = normalizers.BertNormalizer(lowercase=True)
    tok.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    tok.post_processor = processors.TemplateProcessing(single="[CLS] $A [SEP]", special_tokens=[("[CLS]", 2), ("[SEP]", 3)])
    tok.save(path)
    return [w.encode() for w in words]


def make_shard_text(words, n_sentences, shard_index):
    """Raw text of shard `shard_index` (deterministic): sentences of L - 3 words + '.', L from the c5 law."""
    lengths = pkg.synth.lengths_corpus(n_sentences, seed=2026 + shard_index)
    rng = np.random.default_rng(977 + shard_index)
    nw = (lengths - 3).astype(np.int64)  # [CLS] words '.' [SEP]
    pick = rng.integers(0, len(words), size=int(nw.sum()))
    parts, at = [], 0
    for k in nw:
        parts.append(b" ".join(words[i] for i in pick[at:at + k]) + b".")
        at += k
    return b"  ".join(parts), lengths


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0, help="0 = all visible")
    ap.add_argument("--sentences", type=int, default=10_000_000)
    ap.add_argument("--shard", type=int, default=65536)
    ap.add_argument("--base-shards", type=int, default=8)
    ap.add_argument("--inflight", type=int, default=0, help="shards in flight (0 = 2 + gpus / 2)")
    ap.add_argument("--tok-threads", type=int, default=0)
    ap.add_argument("--operands", default="bf16")
    ap.add_argument("--json", action="store_true", help="format the TextWithEmbeddingsMessage payload of every shard as well")
    ap.add_argument("--json-threads", type=int, default=4)
    ap.add_argument("--json-gpu", action="store_true", help="with --json: format on the GPUs (b200_jsongpu_*, one handle per GPU) instead of host threads")
    ap.add_argument("--check", type=int, default=0, help="compare the first N sentences with a direct single-GPU call, in order")
    ap.add_argument("--start-shard", type=int, default=0)
    ap.add_argument("--progress", default="")
    ap.add_argument("--split-threads", type=int, default=0, help="shards split side by side, one thread each (0 = a quarter of the host cores)")
    ap.add_argument("--cold", action="store_true", help="no warm-up pass: the timed region includes every GPU's first-call set-up")
    ap.add_argument("--no-split", action="store_true", help="reuse the sentence ranges of the base shards (tokeniser + GPUs only)")
    args = ap.parse_args()
    import torch
    ngpu = args.gpus or torch.cuda.device_count()
    shape = pkg.weights.MINILM_L6
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tok_path = os.path.join(ROOT, "gpurun_out", "corpus_tokenizer.json")
    words = make_tokenizer(tok_path)
    ingest = pkg.abi.Ingest(tok_path)
    t0 = time.perf_counter()
    base = [make_shard_text(words, args.shard, i) for i in range(args.base_shards)]
    gen_s = time.perf_counter() - t0
    raw_bytes = sum(len(b[0]) for b in base) / len(base)
    n_shards = (args.sentences + args.shard - 1) // args.shard
    inflight = args.inflight or 4 + 2 * ngpu
    cores = len(os.sched_getaffinity(0))
    args.split_threads = args.split_threads or max(2, cores // 4)
    args.tok_threads = args.tok_threads or max(2, cores - ngpu - args.split_threads)  # a worker thread per GPU as well
    H = shape.hidden_size
    corpus = pkg.abi.Corpus(shape, ingest, range(ngpu), tokenizer_threads=args.tok_threads, operands=args.operands)

    result = {"config": f"c5: {args.sentences} sentences of text, MiniLM-L6 {args.operands}, {ngpu} GPU(s), shards of {args.shard}",
              "gpus": ngpu, "host_cores": len(os.sched_getaffinity(0)), "shard_sentences": args.shard, "base_shards": args.base_shards,
              "text_generation_s_untimed": round(gen_s, 2), "raw_text_bytes_per_shard": int(raw_bytes), "inflight_shards": inflight}

    # ---- order / duplication check against a direct single-GPU call -------------------------------------------------------
    if args.check:
        cleaned, begin, end = pkg.abi.split_sentences(base[0][0])
        n = min(args.check, len(begin))
        out = np.zeros((n, H), np.float32)
        corpus.wait(corpus.submit(cleaned, begin[:n], end[:n], out))
        with pkg.Embedder.random_init(shape, seed=1234, device=0, operands=args.operands) as emb:
            ids, tt, cu = ingest.encode_ranges_packed(cleaned, begin[:n], end[:n], shape.max_position_embeddings)
            want = np.concatenate([emb.forward_varlen(ids[cu[a]:cu[b]], tt[cu[a]:cu[b]], cu[a:b + 1] - cu[a], 128)
                                   for a, b in zip(range(0, n, 8192), list(range(8192, n, 8192)) + [n])])
        cos = (out * want).sum(1) / (np.linalg.norm(out, axis=1) * np.linalg.norm(want, axis=1))
        # a permutation or a duplicate would put another sentence's vector in a row: cosine far below 1
        nearest_other = float(np.max((out[:256] @ want[:256].T / np.linalg.norm(out[:256], axis=1)[:, None] / np.linalg.norm(want[:256], axis=1)[None]) - 2 * np.eye(min(256, n))))
        result["check"] = {"sentences": int(n), "cos_min_vs_direct_single_gpu_in_order": float(cos.min()),
                           "max_abs_vs_direct": float(np.abs(out - want).max()), "max_cos_to_a_different_sentence": nearest_other,
                           "lengths_match_law": bool(np.array_equal(np.diff(cu), base[0][1][:n]))}
        assert cos.min() > 0.9999, cos.min()
        corpus.stats(reset=True)

    # ---- the timed ingest --------------------------------------------------------------------------------------------------
    outs = [(pkg.abi.host_alloc if args.json_gpu else np.empty)((args.shard, H), np.float32) for _ in range(inflight)]
    if args.json_gpu:
        import queue
        args.json_threads = ngpu
        jg_free = queue.Queue()
        for g in range(ngpu):  # (formatter of GPU g, its pinned text buffer)
            jg_free.put((pkg.abi.JsonGpu(g), pkg.abi.host_alloc((args.shard * (H * 17 + 48) + int(raw_bytes) * 6 + 4096,), np.uint8)))
    pre_split = [pkg.abi.split_sentences(b[0]) for b in base] if args.no_split else None
    json_pool = ThreadPoolExecutor(max_workers=args.json_threads) if args.json else None
    json_bytes, json_busy, lock = [0], [0.0], threading.Lock()
    checksum = 0.0
    split_s = 0.0
    done_sentences = 0
    pending = []  # (ticket, shard index, slot, cleaned, begin, end, n)
    free_slots = list(range(inflight))
    json_futures = []

    def format_json(slot, cleaned, begin, end, n, shard):
        t = time.perf_counter()
        if args.json_gpu:  # the items on a GPU; the envelope (a few hundred bytes) is the host's
            jg, buf = jg_free.get()
            try:
                doc = jg.sentence_items_host(cleaned, begin, end, outs[slot][:n], out=buf)
            finally:
                jg_free.put((jg, buf))
        else:
            doc = pkg.abi.json_text_with_embeddings(f"doc-{shard}", "http://corpus/synthetic", (cleaned, begin, end), outs[slot][:n],
                                                    "all-MiniLM-L6-v2", 1700000000000 + shard, threads=1, as_array=True)
        with lock:
            json_bytes[0] += len(doc)
            json_busy[0] += time.perf_counter() - t
        return slot

    def retire(block):
        nonlocal checksum, done_sentences
        ticket, shard, slot, cleaned, begin, end, n = pending.pop(0)
        corpus.wait(ticket)
        checksum += float(outs[slot][:n, 0].sum())
        done_sentences += n
        if args.progress and (shard % 8 == 7 or shard == n_shards - 1):
            with open(args.progress, "w") as f:
                f.write(str(shard + 1))
        if json_pool:
            json_futures.append(json_pool.submit(format_json, slot, cleaned, begin, end, n, shard))
        else:
            free_slots.append(slot)

    def reclaim_json(block):
        while json_futures and (block or json_futures[0].done()):
            free_slots.append(json_futures.pop(0).result())
            block = False

    def run(first, last):
        """shards [first, last): split on helper threads (the splitter is called per shard and would serialise on one
        thread), submitted in order, retired in order, at most `inflight` output buffers in use"""
        nonlocal split_s
        split_pool = ThreadPoolExecutor(max_workers=args.split_threads)

        def prepare(shard):
            if pre_split:
                return pre_split[shard % args.base_shards], 0.0
            ts = time.perf_counter()
            r = pkg.abi.split_sentences(base[shard % args.base_shards][0], threads=1)
            return r, time.perf_counter() - ts

        ahead = []  # (shard, future)
        nxt = first
        while nxt < last or ahead or pending or json_futures:
            while nxt < last and len(ahead) < args.split_threads + 1:
                ahead.append((nxt, split_pool.submit(prepare, nxt)))
                nxt += 1
            reclaim_json(False)
            if ahead and free_slots and ahead[0][1].done():
                shard, fut = ahead.pop(0)
                (cleaned, begin, end), dt = fut.result()
                split_s += dt
                slot = free_slots.pop()
                n = min(args.shard, args.sentences - shard * args.shard)
                ticket = corpus.submit(cleaned, begin[:n], end[:n], outs[slot][:n])
                pending.append((ticket, shard, slot, cleaned, begin[:n], end[:n], n))
            elif pending:
                retire(True)
            elif json_futures:
                reclaim_json(True)
            else:
                time.sleep(0.0002)
        split_pool.shutdown()

    # warm-up (untimed): every GPU sees a few micro-batches, so workspaces, function attributes and pinned buffers exist
    warm = min(n_shards - args.start_shard, 2 * ngpu) if not args.cold else 0
    if warm:
        run(args.start_shard, args.start_shard + warm)
        corpus.stats(reset=True)
        checksum, done_sentences, split_s = 0.0, 0, 0.0
        json_bytes[0], json_busy[0] = 0, 0.0
    t_start = time.perf_counter()
    run(args.start_shard, n_shards)
    wall = time.perf_counter() - t_start
    st = corpus.stats()
    corpus.close()
    act = st["per_gpu_active_s"]
    result.update({
        "sentences": done_sentences, "wall_s": round(wall, 3), "sentences_per_s": round(done_sentences / wall),
        "split_in_timed_region": not args.no_split, "split_s_main_thread": round(split_s, 3),
        "tokens": st["tokens"], "micro_batches": st["micro_batches"], "tokenizer_threads": st["tokenizer_threads"],
        "tokenize_busy_s_sum_over_threads": round(st["tokenize_busy_s"], 3), "plan_s": round(st["plan_s"], 3),
        "gpu_worker_s_sum": {"gather": round(st["gather_s"], 3), "blocked_in_wait": round(st["gpu_wait_s"], 3), "scatter": round(st["scatter_s"], 3)},
        "per_gpu_sentences": st["per_gpu_sentences"], "per_gpu_active_s": [round(x, 3) for x in act],
        "imbalance_max_over_mean_tokens": round(max(st["per_gpu_tokens"]) / (sum(st["per_gpu_tokens"]) / len(act)), 4) if act else None,
        "checksum": checksum, "start_shard": args.start_shard, "shards": n_shards, "warmup_shards_untimed": warm,
        "split_threads": args.split_threads,
    })
    if args.json:
        result["json"] = {"bytes": json_bytes[0], "busy_s_sum_over_threads": round(json_busy[0], 3), "threads": args.json_threads,
                          "where": "GPUs (b200_jsongpu_sentence_items_host, one formatter per GPU)" if args.json_gpu else "host threads"}
    print(json.dumps(result), flush=True)


if __name__ == "__main__":
    main()
