"""Helpers for the GPU parity tests: device buffers via torch (plumbing only), bf16 rounding."""
import numpy as np
import torch


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda().contiguous()


def dev_u32(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32)).cuda()


def dev_i32(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)).cuda()


def bf16_round(a):
    """fp32 ndarray rounded to bf16 (RNE) and back: what the library's bf16 copies hold."""
    return torch.from_numpy(np.array(a, dtype=np.float32, copy=True)).bfloat16().float().numpy()


T16 = {"bf16": torch.bfloat16, "f16": torch.float16}


def round16(a, fmt):
    """fp32 ndarray rounded (RNE) to the 16-bit operand format `fmt` ("bf16" / "f16") and back."""
    return torch.from_numpy(np.array(a, dtype=np.float32, copy=True)).to(T16[fmt]).float().numpy()


LO8 = torch.float8_e5m2  # the low half of a (hi, lo) activation: one e5m2 byte per element (csrc/common.cuh)


def dev_lo8(a):
    """fp32 remainders (x - hi) as the library's low-half bytes"""
    return dev(a, LO8)


def nan_lo8(shape):
    """a low-half output buffer filled with NaN bytes (0x7f), so that unwritten elements show"""
    return torch.full(shape, 0x7F, dtype=torch.uint8, device="cuda").view(LO8)


def split_lo8(x, fmt):
    """an fp32 array as a (hi, lo) pair in operand format fmt: (hi + lo as the library reads it back, hi as a device
    tensor, the low-half bytes as a device tensor). hi must be uploaded as it is: rounding hi + lo to 16 bits again can
    land on the neighbour when lo is exactly half an ulp."""
    hi = round16(x, fmt)
    lo = torch.from_numpy(np.ascontiguousarray(x - hi)).to(LO8)
    return hi + lo.float().numpy(), dev(hi, T16[fmt]), lo.cuda().contiguous()


# what hi + lo can be off by, relative to |y|: lo keeps 3 significant bits of a remainder of at most half an ulp of hi
LO8_REL = {"f16": 2.0 ** -14, "bf16": 2.0 ** -11}


def host(t):
    torch.cuda.synchronize()
    return t.float().cpu().numpy()


def padded_to_packed(x, lengths):
    return np.concatenate([x[b, :n] for b, n in enumerate(lengths)], axis=0)


def cu_of(lengths):
    cu = np.zeros(len(lengths) + 1, np.int32)
    np.cumsum(lengths, out=cu[1:])
    return cu
