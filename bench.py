#!/usr/bin/env python
"""bench.py -- sentences/sec of the embedding hot path (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c2|c1|c3|c4|c5]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (mask -> lengths, embed+LN, L encoder layers, masked mean-pool)
over one synthetic tokenised batch per GPU. The default workload is BASELINE.json configs[1]:
all-MiniLM-L6-v2 shape, random init, bf16, batch 4096 x seq 128 per GPU (weak scaling: sentence
batches shard across GPUs with no data-path collective; NCCL is only the barrier / max-over-ranks
plumbing). One JSON line is printed by rank 0:

  value     sentences/s, whole job, inputs already resident in HBM, timed with CUDA events on the
            launching stream (max over ranks)
  e2e       the same metric through the public C-ABI call with pinned HOST buffers: H2D of ids / mask /
            type ids and D2H of the [B,H] f32 embeddings inside the timed region every step
  roofline  the dominant kernel (largest share of the step), from per-kernel CUDA-event timing on its
            stream: algorithmic FLOPs per launch / mean launch duration vs MEASURED_PEAKS.json
  cpu_baseline  the C oracle (a port of the reference's candle CPU path; the reference itself is Rust and
            cannot be built in this image) on the host cores, in the reference's own schedule
            (chunks of 8 sentences, each padded to max_position_embeddings, embedding_generator.rs:146,:75)

--impl reference times that same CPU path as the reference arm (rank 0 only).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

METRIC = "sentences/sec (MiniLM-L6, seq128)"
UNIT = "sentences/s"

WORKLOADS = {
    # name: (shape key, batch per GPU, seq_len, ragged?)
    "c1": ("minilm-l6", 32, 128, False),
    "c2": ("minilm-l6", 4096, 128, False),
    "c3": ("minilm-l6", 2048, 512, True),
    "c4": ("bert-base", 8192, 256, False),
    # SURVEY config c5's length distribution (one shard of the corpus): clip(round(LogNormal(ln 24, 0.6)), 4, 128)
    "c5": ("minilm-l6", 16384, 128, "corpus"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region (pynvml)."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def make_lengths(pkg, batch, seq, ragged, seed):
    if ragged == "corpus":
        return np.minimum(pkg.synth.lengths_corpus(batch, seed=2026 + seed), seq)
    if ragged:
        return pkg.synth.lengths_loguniform(batch, 8, seq, seed=seed)
    return pkg.synth.lengths_dense(batch, seq)


def kernel_work(shape, lengths, kind):
    """Algorithmic (FLOPs, HBM bytes) of ONE launch of a kernel kind for this batch. The kernel's
    roofline is whichever of the two takes longer at the measured peaks."""
    H, I = shape.hidden_size, shape.intermediate_size
    T = float(np.sum(lengths))
    s2 = float(np.sum(np.asarray(lengths, dtype=np.float64) ** 2))
    B = len(lengths)
    if kind == "gemm_qkv":  # x (hi) in, qkv out, bf16
        return 2 * T * 3 * H * H, (2 * H + 6 * H) * T
    if kind == "gemm_attn_out":  # ctx + (hi, lo) residual in, (hi, lo) out
        return 2 * T * H * H, (2 * H + 4 * H + 4 * H) * T
    if kind == "gemm_ffn1":
        return 2 * T * I * H, (2 * H + 2 * I) * T
    if kind == "gemm_ffn2":
        return 2 * T * I * H, (2 * I + 4 * H + 4 * H) * T
    if kind == "ffn_fused":  # (hi, lo) in, (hi, lo) out; the intermediate stays on chip
        return 4 * T * I * H, (4 * H + 4 * H) * T
    if kind == "qkv_attention":  # x (hi) in, ctx out; q, k, v stay on chip
        return 2 * T * 3 * H * H + 4 * H * s2, (2 * H + 2 * H) * T
    if kind == "attention":  # qkv in, ctx out
        return 4 * H * s2, (6 * H + 2 * H) * T
    if kind == "embed_ln":  # fp32 word row in, (hi, lo) bf16 rows out, ids + type ids + mask
        return 0.0, (8 * H + 12) * T
    if kind == "add_ln":  # fp32 GEMM output + (hi, lo) residual in, (hi, lo) out
        return 0.0, 12 * H * T
    if kind == "pool":  # (hi, lo) rows in, f32 [B,H] out
        return 0.0, 4 * H * T + 4 * T + 4 * H * B
    if kind == "lengths":
        return 0.0, 4.0 * B * (T / max(B, 1)) + 8 * B
    raise KeyError(kind)


def cpu_reference_path(pkg, shape, n_sentences, s, schedule, seed=42):
    """The reference's CPU path via the oracle port. schedule "reference": chunks of 8 padded to P
    (embedding_generator.rs:146, :75); "fair": one batch padded to s."""
    from oracle.oracle import Oracle
    w = pkg.weights.init_weights(shape, 1234)
    o = Oracle(shape, w, "f32")
    o.use_all_cores()
    pad = shape.max_position_embeddings if schedule == "reference" else s
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, pkg.synth.lengths_dense(n_sentences, s), pad, seed=seed)
    chunk = 8 if schedule == "reference" else n_sentences
    return o, (ids, tt, mask, chunk)


def run_reference(args, pkg, rank):
    if rank != 0:
        return
    shape_key, batch, seq, ragged = WORKLOADS[args.workload]
    shape = pkg.weights.SHAPES[shape_key]
    n = 8  # one reference chunk per step: a bounded sample of the workload
    o, (ids, tt, mask, chunk) = cpu_reference_path(pkg, shape, n, seq, "reference")
    for _ in range(max(args.warmup, 1)):
        o.generate(ids, tt, mask, chunk=chunk)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        o.generate(ids, tt, mask, chunk=chunk)
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    sample = (f"{n} sentences of {seq} tokens per step, reference schedule: one chunk of 8 padded to "
              f"max_position_embeddings={shape.max_position_embeddings} (embedding_generator.rs:146,:75), f32, "
              "C oracle port of the candle CPU path (the Rust reference cannot be built here)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {shape_key} seq {seq}", "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": o.threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def run_b200(args, pkg, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    shape_key, batch, seq, ragged = WORKLOADS[args.workload]
    if args.batch:
        batch = args.batch
    shape = pkg.weights.SHAPES[shape_key]
    H = shape.hidden_size
    lengths = make_lengths(pkg, batch, seq, ragged, seed=7 + rank)
    ids, mask, tt = pkg.synth.padded_batch(shape.vocab_size, lengths, seq, seed=42 + rank)
    flags = pkg.FLAG_NORMALIZE if args.normalize else 0
    emb = pkg.Embedder.random_init(shape, seed=1234, device=local_rank)

    # ---- leg 1: inputs resident in HBM, CUDA events on the launching stream --------------------
    as_dev = lambda a: torch.from_numpy(a.view(np.int32)).cuda()
    d_ids, d_mask, d_tt = as_dev(ids), as_dev(mask), as_dev(tt)
    d_out = torch.empty((batch, H), dtype=torch.float32, device="cuda")
    stream = torch.cuda.Stream()

    def step_device():
        emb.forward_device(d_ids.data_ptr(), d_mask.data_ptr(), d_tt.data_ptr(), batch, seq, d_out.data_ptr(),
                           flags=flags, stream=stream.cuda_stream)

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step_device()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = emb.kernel_launches()
    with ClockSampler(local_rank) as clocks:
        with torch.cuda.stream(stream):
            ev0.record(stream)
            for _ in range(args.steps):
                step_device()
            ev1.record(stream)
        barrier()
    launches = emb.kernel_launches() - n0
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    value = world * batch * args.steps / (dev_ms * 1e-3)

    # ---- leg 2: public host API, pinned buffers, H2D + D2H inside the timed region ---------------
    nbuf = 2
    pin = [[pkg.abi.host_alloc(a.shape, np.uint32) for a in (ids, mask, tt)] + [pkg.abi.host_alloc((batch, H), np.float32)]
           for _ in range(nbuf)]
    for b in pin:
        b[0][:], b[1][:], b[2][:] = ids, mask, tt

    def e2e_steps(n):
        tickets = [None] * nbuf
        for i in range(n):
            k = i % nbuf
            if tickets[k] is not None:
                emb.wait(tickets[k])  # result of step i - nbuf is on the host
            tickets[k] = emb.submit(pin[k][0], pin[k][1], pin[k][2], pin[k][3], flags=flags)
        for t in tickets:
            if t is not None:
                emb.wait(t)

    e2e_steps(args.warmup)
    barrier()
    t0 = time.perf_counter()
    e2e_steps(args.steps)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = world * batch * args.steps / e2e_s
    e2e_check = float(np.abs(pin[0][3] - d_out.cpu().numpy()).max())

    # ---- per-kernel timing (outside the timed regions): CUDA events around every launch ----------
    emb.profile(True)
    emb.profile_read(reset=True)
    prof_steps = 3
    for _ in range(prof_steps):
        emb.forward_device(d_ids.data_ptr(), d_mask.data_ptr(), d_tt.data_ptr(), batch, seq, d_out.data_ptr(), flags=flags)
    stats = emb.profile_read(reset=True)
    emb.profile(False)
    pk = peaks()
    kernels, total_ms = {}, sum(v["total_ms"] for v in stats.values()) / prof_steps
    per_layer = {"gemm_qkv", "attention", "gemm_attn_out", "gemm_ffn1", "gemm_ffn2", "ffn_fused", "qkv_attention"}
    for kind, v in stats.items():
        if v["launches"] == 0:
            continue
        ms = v["total_ms"] / v["launches"]
        # work of one full pass of this kind over the batch x how many passes a step makes (calls larger than the
        # workspace are split into several launches per pass; their times add up)
        passes = (2 if kind == "add_ln" else 1) * (shape.num_hidden_layers if kind in per_layer or kind == "add_ln" else 1)
        flops, nbytes = kernel_work(shape, lengths, kind)
        step_ms = v["total_ms"] / prof_steps
        t_tensor = flops * passes / (pk["tf_sustained"] * 1e12) * 1e3  # ms this kind needs at the tensor peak
        t_hbm = nbytes * passes / (pk["hbm_gbs"] * 1e9) * 1e3           # ms at the HBM peak
        if t_tensor >= t_hbm:
            ach, peak, unit, bound = flops * passes / (step_ms * 1e-3) / 1e12, pk["tf_sustained"], "TFLOP/s", "tensor"
        else:
            ach, peak, unit, bound = nbytes * passes / (step_ms * 1e-3) / 1e9, pk["hbm_gbs"], "GB/s", "hbm"
        kernels[kind] = {"ms_per_launch": round(ms, 4), "launches_per_step": v["launches"] // prof_steps,
                         "share_of_step": round(step_ms / total_ms, 4), "bound": bound,
                         "achieved": round(ach, 2), "peak": peak, "unit": unit, "frac": round(ach / peak, 4)}
    top = max(kernels, key=lambda k: kernels[k]["share_of_step"])
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")  # dram bytes per launch from the committed ncu capture
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(f"{args.workload}:{top}")
    roofline = {"kernel": top, "bound": kernels[top]["bound"], "achieved": kernels[top]["achieved"],
                "peak": kernels[top]["peak"], "unit": kernels[top]["unit"], "frac": kernels[top]["frac"],
                "traffic": traffic, "peak_source": pk["source"] + ", sustained figure (kernel timed inside a long step)",
                "ms_per_launch": kernels[top]["ms_per_launch"], "share_of_step": kernels[top]["share_of_step"]}
    uniq, cnt = np.unique(lengths, return_counts=True)
    flops_per_step = float(sum(shape.flops_per_sentence(int(s)) * int(c) for s, c in zip(uniq, cnt)))
    model_tflops = flops_per_step * args.steps / (dev_ms * 1e-3) / 1e12  # per GPU

    # ---- CPU baseline (rank 0, N = 1 only) ---------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n = 16
        o, (ci, ct, cm, chunk) = cpu_reference_path(pkg, shape, n, seq, "reference")
        o.generate(ci[:8], ct[:8], cm[:8], chunk=chunk)
        t0 = time.perf_counter()
        ref_out = o.generate(ci, ct, cm, chunk=chunk)
        dt = time.perf_counter() - t0
        o2, (fi, ft, fm, fchunk) = cpu_reference_path(pkg, shape, 32, seq, "fair")
        t0 = time.perf_counter()
        o2.generate(fi, ft, fm, chunk=fchunk)
        dt_fair = time.perf_counter() - t0
        cpu = {"value": n / dt, "unit": UNIT, "cores": o.threads, "kind": "port",
               "sample": (f"{n} sentences of {seq} tokens, reference schedule (chunks of 8, padded to "
                          f"P={shape.max_position_embeddings}, f32), C oracle port of the candle CPU path"),
               "fair_pad_to_seq_b32": 32 / dt_fair}
        del ref_out

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {shape_key} random-init bf16, batch {batch} x seq {seq} per GPU"
                                   + (" corpus lengths clip(LogNormal(ln 24, 0.6), 4, 128)" if ragged == "corpus"
                                      else " ragged log-uniform 8..seq packed varlen" if ragged else " dense"),
                       "tokens_per_step_per_gpu": int(np.sum(lengths)), "normalize": bool(args.normalize),
                       "parallelism": f"replica per GPU x{world}, sentence batches sharded, no data-path collective",
                       "l2": "activations per step (>= 4 GB at c2) exceed the 126 MB L2; no explicit flush"},
            "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(3 * ids.nbytes),
                    "d2h_bytes_per_step": int(batch * H * 4), "ms_per_step": 1e3 * e2e_s / args.steps,
                    "api": "b200_embed_submit/b200_embed_wait, 2 requests in flight, pinned host buffers",
                    "max_abs_vs_device_leg": e2e_check},
            "gpu_launches": int(launches),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "kernels": kernels,
            "model_tflops_per_gpu": round(model_tflops, 1),
            "model_frac_of_sustained_bf16_peak": round(model_tflops / pk["tf_sustained"], 4),
        }
        print(json.dumps(line), flush=True)
    for b in pin:
        for a in b:
            pkg.abi.host_free(a)
    emb.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="override sentences per GPU per step")
    ap.add_argument("--normalize", action="store_true", help="append the L2 normalisation (north_star option)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    pkg = load_package()
    if args.impl == "reference":
        run_reference(args, pkg, rank)
        return
    if world == 1 and args.gpus > 1:
        raise SystemExit("bench.py: --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    run_b200(args, pkg, rank, world, local_rank)


if __name__ == "__main__":
    main()
