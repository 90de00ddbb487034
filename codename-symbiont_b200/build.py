"""nvcc recipe for csrc/ -> libb200_embed.so (sm_100a only, built in-tree so it travels to the GPU box)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200_embed.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = ["engine.cu", "fp32_ref.cu", "weights_io.cpp", "json_out.cpp", "ingest.cpp", "search.cu", "pipeline.cpp", "corpus.cpp", "json_gpu.cu"]
# The 16-bit kernels are compiled once per operand format (csrc/fmt.h): as they are for bf16, with -DB200_OPERAND_F16 for
# IEEE fp16; each build lives in its own namespace and the engine picks per handle.
FORMAT_SOURCES = ["gemm_tc.cu", "ffn_fused.cu", "attention_tc.cu", "qkv_attention.cu", "elementwise.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
              "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function", "--expt-relaxed-constexpr"]


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def build_library(verbose: bool = False, force: bool = False, suffix: str = "", extra_flags=()) -> str:
    """suffix / extra_flags: an experimental variant next to the shipped library (libb200_embed<suffix>.so, objects in
    build<suffix>/), loaded with B200_EMBED_LIB=... for same-box A/B timing (scripts/k23_ab.py)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    LIB = os.path.join(HERE, f"libb200_embed{suffix}.so")
    OBJ_DIR = os.path.join(HERE, "build" + suffix)
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "b200_embed.h"))
    jobs = []
    objs = []
    units = [(src, "", []) for src in SOURCES] + [(src, "", []) for src in FORMAT_SOURCES] + \
            [(src, "_f16", ["-DB200_OPERAND_F16"]) for src in FORMAT_SOURCES]
    for src, suffix, defs in units:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ_DIR, os.path.splitext(src)[0] + suffix + ".o")
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            jobs.append([nvcc, *NVCC_FLAGS, *defs, *extra_flags, "-c", s, "-o", o])
    jobs.sort(key=lambda j: 0 if "gemm_tc" in j[-3] else 1)  # the longest compiles first

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed:\n{r.stdout}\n{r.stderr}")
        return r

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
        list(ex.map(run, jobs))
    if jobs or force or not _newer(LIB, objs):
        run([nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"])
    return LIB


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--suffix", default="")
    ap.add_argument("--flags", default="", help="extra nvcc flags of a variant build, e.g. '-DQA_POLY_PAIRS=0'")
    a = ap.parse_args()
    print(build_library(verbose=True, force=a.force, suffix=a.suffix, extra_flags=a.flags.split()))
