"""B200-native sentence-embedding path behind Symbiont's `EmbeddingGenerator`.

Host side (Python, because the reference's Rust toolchain is absent from this image):
  abi                  ctypes binding of libb200_embed.so (include/b200_embed.h)
  embedding_generator  mirror of the reference's `EmbeddingGenerator` interface
  weights / synth      model shapes, random-init spec, safetensors IO, synthetic token batches
  sharding             micro-batching and dealing sentence batches to the GPUs of one box
  build                the nvcc recipe for csrc/ (sm_100a)
The compute lives in csrc/ (hand-written CUDA behind a C ABI). There is no CPU fallback.
"""
from . import weights, synth, build, abi, sharding  # noqa: F401
from .abi import (Embedder, B200EmbedError, FLAG_NORMALIZE, FLAG_FP32)  # noqa: F401
from .embedding_generator import EmbeddingGenerator  # noqa: F401
