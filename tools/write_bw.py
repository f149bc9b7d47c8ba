"""Pure-write HBM bandwidth on this GPU (cudaMemset / fill kernel), the ceiling of write-only kernels such as the
stand-alone ScatterConnection (3.1 MB written per observation)."""
import torch
n = 3306700800 // 4
x = torch.empty(n, dtype=torch.float32, device='cuda')
for name, fn in (('memset (zero_)', lambda: x.zero_()), ('fill kernel (fill_ 1.0)', lambda: x.fill_(1.0))):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print('%-26s %.3f ms  %.0f GB/s' % (name, ms, n * 4 / ms / 1e6))
y = torch.empty_like(x)
for _ in range(3):
    y.copy_(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    y.copy_(x)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print('%-26s %.3f ms  %.0f GB/s (read + write)' % ('copy', ms, 2 * n * 4 / ms / 1e6))
