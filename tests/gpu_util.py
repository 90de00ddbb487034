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


def host(t):
    torch.cuda.synchronize()
    return t.float().cpu().numpy()


def padded_to_packed(x, lengths):
    return np.concatenate([x[b, :n] for b, n in enumerate(lengths)], axis=0)


def cu_of(lengths):
    cu = np.zeros(len(lengths) + 1, np.int32)
    np.cumsum(lengths, out=cu[1:])
    return cu
