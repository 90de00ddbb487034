"""Model shapes, the deterministic random-init spec, and a minimal safetensors reader/writer.

The reference loads HuggingFace BERT weights from `model.safetensors` at F32
(/root/reference/services/preprocessing_service/src/embedding_generator.rs:106-124) using the
state-dict names candle's `BertModel::load` consumes. There is no
network here, so benchmarks and tests use random-init weights of the named architecture.
The init spec below is implemented twice, bit-for-bit in integer space: here in numpy and
natively in csrc/weights_init.cpp (`b200_embed_create` with a NULL path):

    key      = fnv1a64(tensor_name) XOR (seed * 0x9E3779B97F4A7C15 mod 2^64)
    a, b     = splitmix64(key + 2i), splitmix64(key + 2i + 1)          (element i)
    u1, u2   = ((a >> 11) + 1) * 2^-53 ,  (b >> 11) * 2^-53
    z        = sqrt(-2 ln u1) * cos(2 pi u2)                           (double)
    value    = float32(mean + 0.02 * z)       mean = 1 for LayerNorm.weight, else 0
"""
from __future__ import annotations

import json
import struct
from dataclasses import dataclass, asdict

import numpy as np

ACT_GELU_ERF, ACT_GELU_TANH, ACT_RELU = 0, 1, 2
_ACT_NAMES = {"gelu": ACT_GELU_ERF, "gelu_new": ACT_GELU_TANH, "gelu_pytorch_tanh": ACT_GELU_TANH,
              "gelu_approximate": ACT_GELU_TANH, "relu": ACT_RELU}


@dataclass(frozen=True)
class BertShape:
    """The `config.json` fields candle's `bert::Config` reads (embedding_generator.rs:70-71)."""
    vocab_size: int
    hidden_size: int
    num_hidden_layers: int
    num_attention_heads: int
    intermediate_size: int
    max_position_embeddings: int
    type_vocab_size: int
    hidden_act: int = ACT_GELU_ERF
    layer_norm_eps: float = 1e-12

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    def flops_per_sentence(self, s: int) -> float:
        """F(s) = L*(24*H^2*s + 4*H*s^2)  (GEMM + attention MACs x2, I = 4H)."""
        L, H, I = self.num_hidden_layers, self.hidden_size, self.intermediate_size
        return float(L * ((8 * H * H + 4 * H * I) * s + 4 * H * s * s))

    @staticmethod
    def from_config_json(path: str) -> "BertShape":
        with open(path) as f:
            c = json.load(f)
        act = c.get("hidden_act", "gelu")
        if act not in _ACT_NAMES:
            raise ValueError(f"unsupported hidden_act {act!r}")
        return BertShape(c["vocab_size"], c["hidden_size"], c["num_hidden_layers"], c["num_attention_heads"],
                         c["intermediate_size"], c["max_position_embeddings"], c.get("type_vocab_size", 2),
                         _ACT_NAMES[act], float(c.get("layer_norm_eps", 1e-12)))

    def to_dict(self):
        return asdict(self)


# BASELINE.json shapes (SURVEY.md section 8 header).
MINILM_L6 = BertShape(30522, 384, 6, 12, 1536, 512, 2, ACT_GELU_ERF, 1e-12)
BERT_BASE = BertShape(30522, 768, 12, 12, 3072, 512, 2, ACT_GELU_ERF, 1e-12)
XLMR_MPNET_BASE = BertShape(250002, 768, 12, 12, 3072, 514, 1, ACT_GELU_ERF, 1e-5)  # production model, main.rs:305
TINY = BertShape(512, 128, 2, 4, 512, 64, 2, ACT_GELU_ERF, 1e-12)                   # golden-fixture shape
SHAPES = {"minilm-l6": MINILM_L6, "bert-base": BERT_BASE, "xlmr-mpnet-base": XLMR_MPNET_BASE, "tiny": TINY}


def tensor_specs(shape: BertShape):
    """(name, dims) in the HF BERT state-dict layout candle consumes."""
    H, I = shape.hidden_size, shape.intermediate_size
    out = [("embeddings.word_embeddings.weight", (shape.vocab_size, H)),
           ("embeddings.position_embeddings.weight", (shape.max_position_embeddings, H)),
           ("embeddings.token_type_embeddings.weight", (shape.type_vocab_size, H)),
           ("embeddings.LayerNorm.weight", (H,)), ("embeddings.LayerNorm.bias", (H,))]
    for i in range(shape.num_hidden_layers):
        p = f"encoder.layer.{i}."
        out += [(p + "attention.self.query.weight", (H, H)), (p + "attention.self.query.bias", (H,)),
                (p + "attention.self.key.weight", (H, H)), (p + "attention.self.key.bias", (H,)),
                (p + "attention.self.value.weight", (H, H)), (p + "attention.self.value.bias", (H,)),
                (p + "attention.output.dense.weight", (H, H)), (p + "attention.output.dense.bias", (H,)),
                (p + "attention.output.LayerNorm.weight", (H,)), (p + "attention.output.LayerNorm.bias", (H,)),
                (p + "intermediate.dense.weight", (I, H)), (p + "intermediate.dense.bias", (I,)),
                (p + "output.dense.weight", (H, I)), (p + "output.dense.bias", (H,)),
                (p + "output.LayerNorm.weight", (H,)), (p + "output.LayerNorm.bias", (H,))]
    return out


_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _fnv1a64(name: str) -> int:
    h = 0xCBF29CE484222325
    for b in name.encode():
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def init_tensor(name: str, dims, seed: int) -> np.ndarray:
    n = int(np.prod(dims))
    key = np.uint64(_fnv1a64(name) ^ ((seed * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF))
    out = np.empty(n, dtype=np.float32)
    mean = 1.0 if name.endswith("LayerNorm.weight") else 0.0
    step = 1 << 22
    for lo in range(0, n, step):
        i = np.arange(lo, min(n, lo + step), dtype=np.uint64)
        with np.errstate(over="ignore"):
            a = _splitmix64(key + np.uint64(2) * i)
            b = _splitmix64(key + np.uint64(2) * i + np.uint64(1))
        u1 = ((a >> np.uint64(11)).astype(np.float64) + 1.0) * (2.0 ** -53)
        u2 = (b >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)
        z = np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
        out[lo:lo + len(i)] = (mean + 0.02 * z).astype(np.float32)
    return out.reshape(dims)


def init_weights(shape: BertShape, seed: int = 1234) -> dict:
    return {name: init_tensor(name, dims, seed) for name, dims in tensor_specs(shape)}


# ---- safetensors: u64 LE header length, JSON header, raw little-endian tensor bytes ----
_ST_DTYPES = {"F32": np.float32, "F16": np.float16, "F64": np.float64, "I64": np.int64, "I32": np.int32}


def to_bf16_bits(a: np.ndarray) -> np.ndarray:
    """f32 -> bf16 bit patterns (uint16), round to nearest even."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    return ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)


def save_safetensors(path: str, tensors: dict, metadata: dict | None = None, store_as: str | None = None) -> None:
    """store_as: None keeps each array's dtype (F32 / F16 / F64); "F16" / "BF16" converts every tensor on the way out
    (checkpoints are published in all three)."""
    header, off = {}, 0
    if metadata:
        header["__metadata__"] = {k: str(v) for k, v in metadata.items()}
    blobs = []
    for name, arr in tensors.items():
        arr = np.ascontiguousarray(arr)
        if store_as == "BF16":
            dt, raw = "BF16", to_bf16_bits(arr).tobytes()
        else:
            if store_as == "F16":
                arr = arr.astype(np.float16)
            dt = {np.dtype(np.float32): "F32", np.dtype(np.float16): "F16", np.dtype(np.float64): "F64"}[arr.dtype]
            raw = arr.tobytes()
        nb = len(raw)
        header[name] = {"dtype": dt, "shape": list(arr.shape), "data_offsets": [off, off + nb]}
        blobs.append(raw)
        off += nb
    hj = json.dumps(header, separators=(",", ":")).encode()
    hj += b" " * ((8 - len(hj) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hj)))
        f.write(hj)
        for b in blobs:
            f.write(b)


def load_safetensors(path: str) -> dict:
    with open(path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        header = json.loads(f.read(n))
        base = 8 + n
        out = {}
        for name, info in header.items():
            if name == "__metadata__":
                continue
            lo, hi = info["data_offsets"]
            f.seek(base + lo)
            raw = f.read(hi - lo)
            if info["dtype"] == "BF16":
                u = np.frombuffer(raw, dtype=np.uint16).astype(np.uint32) << 16
                arr = u.view(np.float32)
            else:
                arr = np.frombuffer(raw, dtype=_ST_DTYPES[info["dtype"]])
            out[name] = arr.reshape(info["shape"]).astype(np.float32, copy=False)
    return out
