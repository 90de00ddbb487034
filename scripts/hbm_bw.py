import torch
def t(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
N = 1 << 30  # 1 Gi bf16 = 2 GiB
a = torch.empty(N, dtype=torch.bfloat16, device="cuda"); b = torch.empty(N, dtype=torch.bfloat16, device="cuda")
ms = t(lambda: a.zero_()); print(f"write only (zero_ 2 GiB): {ms:.3f} ms  {2*N/ms/1e6:.0f} GB/s")
ms = t(lambda: a.fill_(1.5)); print(f"write only (fill_ 2 GiB): {ms:.3f} ms  {2*N/ms/1e6:.0f} GB/s")
ms = t(lambda: b.copy_(a)); print(f"copy 2 GiB -> 2 GiB: {ms:.3f} ms  {4*N/ms/1e6:.0f} GB/s (read+write)")
ms = t(lambda: a.sum()); print(f"read only (sum 2 GiB): {ms:.3f} ms  {2*N/ms/1e6:.0f} GB/s")
c = torch.empty(N // 4, dtype=torch.bfloat16, device="cuda")
ms = t(lambda: torch.cat([c, c, c], out=a[: 3 * (N // 4)])); print(f"read 0.5 GiB x3 (L2?) write 1.5 GiB: {ms:.3f} ms  write {1.5*N/ms/1e6*1:.0f} GB/s")
