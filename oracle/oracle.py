"""ctypes binding of oracle/bert_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module (see the header of bert_oracle.c for the parity status: the reference
itself pins nothing for this path; this repo pins the oracle against HuggingFace
transformers via oracle/make_golden.py -> tests/golden/).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class OracleConfig(C.Structure):
    _fields_ = [("vocab_size", C.c_int32), ("hidden_size", C.c_int32), ("num_layers", C.c_int32),
                ("num_heads", C.c_int32), ("intermediate_size", C.c_int32),
                ("max_position_embeddings", C.c_int32), ("type_vocab_size", C.c_int32),
                ("hidden_act", C.c_int32), ("layer_norm_eps", C.c_float)]


_LAYER_FIELDS = ["q_w", "q_b", "k_w", "k_b", "v_w", "v_b", "ao_w", "ao_b", "ao_ln_g", "ao_ln_b",
                 "i_w", "i_b", "o_w", "o_b", "o_ln_g", "o_ln_b"]
_LAYER_NAMES = ["attention.self.query.weight", "attention.self.query.bias", "attention.self.key.weight",
                "attention.self.key.bias", "attention.self.value.weight", "attention.self.value.bias",
                "attention.output.dense.weight", "attention.output.dense.bias",
                "attention.output.LayerNorm.weight", "attention.output.LayerNorm.bias",
                "intermediate.dense.weight", "intermediate.dense.bias", "output.dense.weight",
                "output.dense.bias", "output.LayerNorm.weight", "output.LayerNorm.bias"]


class OracleLayer(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _LAYER_FIELDS]


class OracleWeights(C.Structure):
    _fields_ = [("word_emb", C.c_void_p), ("pos_emb", C.c_void_p), ("type_emb", C.c_void_p),
                ("emb_ln_g", C.c_void_p), ("emb_ln_b", C.c_void_p), ("layers", C.POINTER(OracleLayer))]


def build(force: bool = False) -> None:
    """Compile liboracle_f32.so / liboracle_f64.so with gcc (oracle/Makefile)."""
    if force or not all(os.path.exists(os.path.join(_HERE, f)) for f in ("liboracle_f32.so", "liboracle_f64.so")):
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def build_native() -> str:
    """liboracle_f32_native.so: the same source built for THIS host (-march=native, contraction allowed), used only for
    the timed CPU-baseline legs of bench.py so that the baseline is not artificially slow. Rebuilt on every call: a
    library built for another machine's instruction set must not travel."""
    out = os.path.join(_HERE, "_native", "liboracle_f32_native.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.run(["gcc", "-O3", "-march=native", "-fopenmp", "-fPIC", "-shared", "-fno-fast-math", "-ffp-contract=fast",
                    "-o", out, os.path.join(_HERE, "bert_oracle.c"), "-lm"], check=True)
    return out


class Oracle:
    """The reference path on the CPU. `precision` is "f32" (candle's DTYPE) or "f64"."""

    _native_path = None

    def __init__(self, shape, weights: dict, precision: str = "f32", native: bool = False):
        build()
        if native and precision == "f32":
            if Oracle._native_path is None:
                Oracle._native_path = build_native()
            self.lib = C.CDLL(Oracle._native_path)
        else:
            self.lib = C.CDLL(os.path.join(_HERE, f"liboracle_{precision}.so"))
        self.real = np.float32 if precision == "f32" else np.float64
        assert self.lib.oracle_real_bytes() == np.dtype(self.real).itemsize
        self.shape = shape
        self.cfg = OracleConfig(shape.vocab_size, shape.hidden_size, shape.num_hidden_layers,
                                shape.num_attention_heads, shape.intermediate_size,
                                shape.max_position_embeddings, shape.type_vocab_size, shape.hidden_act,
                                shape.layer_norm_eps)
        self._keep = {k: _f32(v) for k, v in weights.items()}
        p = lambda name: self._keep[name].ctypes.data_as(C.c_void_p)
        self._layers = (OracleLayer * shape.num_hidden_layers)()
        for i in range(shape.num_hidden_layers):
            for f, n in zip(_LAYER_FIELDS, _LAYER_NAMES):
                setattr(self._layers[i], f, p(f"encoder.layer.{i}.{n}"))
        self.w = OracleWeights(p("embeddings.word_embeddings.weight"), p("embeddings.position_embeddings.weight"),
                               p("embeddings.token_type_embeddings.weight"), p("embeddings.LayerNorm.weight"),
                               p("embeddings.LayerNorm.bias"), self._layers)
        self.lib.oracle_linear.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                           C.c_void_p]
        self.lib.oracle_activation.argtypes = [C.c_void_p, C.c_int64, C.c_int]
        self.lib.oracle_add_layer_norm.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p,
                                                   C.c_void_p, C.c_float]
        self.lib.oracle_layer_norm.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_float]

    @property
    def threads(self) -> int:
        return int(self.lib.oracle_num_threads())

    def use_all_cores(self) -> int:
        """torchrun exports OMP_NUM_THREADS=1; the CPU baseline is meant to use every host core."""
        n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.lib.oracle_set_num_threads(n)
        return self.threads

    def layer(self, i: int, name: str) -> np.ndarray:
        return self._keep[f"encoder.layer.{i}.{name}"]

    # -- whole path -------------------------------------------------------------------------
    def bert_forward(self, ids, type_ids, mask) -> np.ndarray:
        """BertModel::forward -> last hidden state [B,S,H] (embedding_generator.rs:198)."""
        ids, type_ids, mask = _u32(ids), _u32(type_ids), _u32(mask)
        B, S = ids.shape
        out = np.empty((B, S, self.shape.hidden_size), dtype=self.real)
        rc = self.lib.oracle_bert_forward(C.byref(self.cfg), C.byref(self.w), ids.ctypes, type_ids.ctypes,
                                          mask.ctypes, B, S, out.ctypes)
        if rc:
            raise ValueError(f"oracle_bert_forward failed: {rc}")
        return out

    def mean_pool(self, hidden, mask) -> np.ndarray:
        """embedding_generator.rs:201-207."""
        hidden = np.ascontiguousarray(hidden, dtype=self.real)
        mask = _u32(mask)
        B, S, H = hidden.shape
        out = np.empty((B, H), dtype=np.float32)
        self.lib.oracle_mean_pool(hidden.ctypes, mask.ctypes, B, S, H, out.ctypes)
        return out

    def l2_normalize(self, emb) -> np.ndarray:
        out = _f32(emb).copy()
        self.lib.oracle_l2_normalize(out.ctypes, out.shape[0], out.shape[1])
        return out

    def generate(self, ids, type_ids, mask, chunk: int = 8, normalize: bool = False) -> np.ndarray:
        """generate_sentence_embeddings after tokenisation: [N,P] u32 x3 -> [N,H] f32."""
        ids, type_ids, mask = _u32(ids), _u32(type_ids), _u32(mask)
        N, P = ids.shape
        out = np.empty((N, self.shape.hidden_size), dtype=np.float32)
        rc = self.lib.oracle_generate(C.byref(self.cfg), C.byref(self.w), ids.ctypes, type_ids.ctypes, mask.ctypes,
                                      N, P, chunk, int(normalize), out.ctypes)
        if rc:
            raise ValueError(f"oracle_generate failed: {rc}")
        return out

    # -- single ops (kernel-level parity tests) ---------------------------------------------
    def embeddings(self, ids, type_ids) -> np.ndarray:
        ids, type_ids = _u32(ids), _u32(type_ids)
        B, S = ids.shape
        out = np.empty((B, S, self.shape.hidden_size), dtype=self.real)
        self.lib.oracle_embeddings(C.byref(self.cfg), C.byref(self.w), ids.ctypes, type_ids.ctypes, B, S, out.ctypes)
        return out

    def linear(self, x, w, b) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=self.real)
        w, b = _f32(w), _f32(b)
        M, K = x.shape
        N = w.shape[0]
        y = np.empty((M, N), dtype=self.real)
        self.lib.oracle_linear(x.ctypes.data, M, K, w.ctypes.data, b.ctypes.data, N, y.ctypes.data)
        return y

    def activation(self, x, act: int = 0) -> np.ndarray:
        y = np.ascontiguousarray(x, dtype=self.real).copy()
        self.lib.oracle_activation(y.ctypes.data, y.size, act)
        return y

    def add_layer_norm(self, y, residual, g, b, eps) -> np.ndarray:
        y = np.ascontiguousarray(y, dtype=self.real).copy()
        r = np.ascontiguousarray(residual, dtype=self.real)
        g, b = _f32(g), _f32(b)
        self.lib.oracle_add_layer_norm(y.ctypes.data, r.ctypes.data, y.shape[0], y.shape[1], g.ctypes.data,
                                       b.ctypes.data, eps)
        return y

    def attention(self, q, k, v, mask, heads: int) -> np.ndarray:
        q = np.ascontiguousarray(q, dtype=self.real)
        k = np.ascontiguousarray(k, dtype=self.real)
        v = np.ascontiguousarray(v, dtype=self.real)
        mask = _u32(mask)
        B, S, H = q.shape
        ctx = np.empty_like(q)
        self.lib.oracle_attention(q.ctypes, k.ctypes, v.ctypes, mask.ctypes, B, S, H, heads, ctx.ctypes)
        return ctx
