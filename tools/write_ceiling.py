"""Write-only HBM ceiling on this GPU: torch fill / cudaMemset of an obs-sized buffer (4.93 GB)."""
import torch
n = (1 << 20) * 4 * 49 * 6
x = torch.empty(n, dtype=torch.float32, device='cuda')
def t(f, reps=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for name, f in [('zero_ (memset)', lambda: x.zero_()), ('fill_(1.0)', lambda: x.fill_(1.0))]:
    ms = t(f)
    print(f'{name}: {ms:.3f} ms  {n * 4 / ms / 1e6:.0f} GB/s')
y = torch.empty(n // 2, dtype=torch.float32, device='cuda'); z = torch.empty_like(y)
ms = t(lambda: z.copy_(y)); print(f'copy 2x{n*2/1e9:.2f} GB: {ms:.3f} ms {n * 4 / ms / 1e6:.0f} GB/s (read+write)')
